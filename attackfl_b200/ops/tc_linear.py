"""tcgen05 linear layers (``csrc/gemm_tcgen05.cu``): forward, input gradient and weight gradient of ``nn.Linear`` on
fp32 activations read as TF32 tensor-core operands, with the bias / ReLU / dropout / mask epilogues fused.

Used by the native HAR trainer (``ops.har``); ``TcLinear`` wraps the three kernels as one autograd function so the
PyTorch module path can use them as a drop-in for ``F.linear`` on CUDA tensors.  Reference counterpart: the
``nn.Linear`` calls inside ``nn.TransformerEncoderLayer`` (``/root/reference/src/Model.py:435-458``)."""
from __future__ import annotations

import ctypes
from typing import Dict, Optional, Tuple

import torch

from .native import check, cuda_lib, ptr, stream_ptr

c_i, c_f, c_i64, c_u64, c_u32 = ctypes.c_int, ctypes.c_float, ctypes.c_int64, ctypes.c_uint64, ctypes.c_uint32
_ws: Dict[Tuple[int, int, int], torch.Tensor] = {}


def supported(n: int, k: int) -> bool:
    return 32 <= n <= 256 and 32 <= k <= 256 and n % 32 == 0 and k % 32 == 0


def linear_fwd(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor] = None, relu: bool = False,
               drop_p: float = 0.0, seed: int = 0, rng: Tuple[int, int, int] = (0, 0, 0),
               out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``dropout(relu?(x @ w.T + bias))`` for ``x [M, K]``, ``w [N, K]``; ``rng = (client, step, rng_id)``."""
    m, k = x.shape
    n = w.shape[0]
    lib = cuda_lib()
    out = torch.empty(m, n, dtype=torch.float32, device=x.device) if out is None else out
    rc = lib.afl_tc_linear_fwd(ptr(x), c_i64(m), c_i(k), ptr(w), c_i(n), ptr(bias), c_i(1 if relu else 0), c_f(drop_p),
                               c_u64(seed), c_u32(rng[0]), c_u32(rng[1]), c_u32(rng[2]), ptr(out), stream_ptr(x))
    check(lib, rc, "afl_tc_linear_fwd")
    return out


def linear_ln_fwd(x: torch.Tensor, w: torch.Tensor, bias: torch.Tensor, resid: torch.Tensor, gamma: torch.Tensor,
                  beta: torch.Tensor, y: torch.Tensor, xhat: torch.Tensor, rstd: torch.Tensor, drop_p: float = 0.0,
                  seed: int = 0, rng: Tuple[int, int, int] = (0, 0, 0)) -> torch.Tensor:
    """``y = LayerNorm(resid + dropout(x @ w.T + bias)) * gamma + beta`` for ``w [64, K]`` in one kernel (the normalisation
    runs in the GEMM epilogue); ``xhat``, ``rstd`` are saved for the backward pass."""
    m, k = x.shape
    lib = cuda_lib()
    rc = lib.afl_tc_linear_ln_fwd(ptr(x), c_i64(m), c_i(k), ptr(w), ptr(bias), ptr(resid), c_f(drop_p), c_u64(seed),
                                  c_u32(rng[0]), c_u32(rng[1]), c_u32(rng[2]), ptr(gamma), ptr(beta), ptr(y), ptr(xhat),
                                  ptr(rstd), stream_ptr(x))
    check(lib, rc, "afl_tc_linear_ln_fwd")
    return y


def linear_dx(dy: torch.Tensor, w: torch.Tensor, mask_src: Optional[torch.Tensor] = None, mask_scale: float = 1.0,
              out: Optional[torch.Tensor] = None, accumulate: bool = False) -> torch.Tensor:
    """``(dy @ w) * (mask_src > 0 ? mask_scale : 0)`` for ``dy [M, N]``, ``w [N, K]``."""
    m, n = dy.shape
    k = w.shape[1]
    lib = cuda_lib()
    out = torch.empty(m, k, dtype=torch.float32, device=dy.device) if out is None else out
    rc = lib.afl_tc_linear_dx(ptr(dy), c_i64(m), c_i(n), ptr(w), c_i(k), ptr(mask_src), c_f(mask_scale),
                              c_i(1 if accumulate else 0), ptr(out), stream_ptr(dy))
    check(lib, rc, "afl_tc_linear_dx")
    return out


def _workspace(n: int, k: int, device) -> torch.Tensor:
    key = (n, k, torch.device(device).index or 0)
    if key not in _ws:
        lib = cuda_lib()
        need = c_i64(0)
        check(lib, lib.afl_tc_linear_dw_workspace(c_i(n), c_i(k), ctypes.byref(need)), "afl_tc_linear_dw_workspace")
        _ws[key] = torch.empty(need.value, dtype=torch.float32, device=device)
    return _ws[key]


def linear_dw(dy: torch.Tensor, x: torch.Tensor, dw: Optional[torch.Tensor] = None, db: Optional[torch.Tensor] = None,
              want_db: bool = True, accumulate: bool = False):
    """``dw = dy.T @ x`` (``[N, K]``) and ``db = dy.sum(0)``."""
    m, n = dy.shape
    k = x.shape[1]
    lib = cuda_lib()
    dw = torch.empty(n, k, dtype=torch.float32, device=dy.device) if dw is None else dw
    if want_db and db is None:
        db = torch.empty(n, dtype=torch.float32, device=dy.device)
    rc = lib.afl_tc_linear_dw(ptr(dy), ptr(x), c_i64(m), c_i(n), c_i(k), ptr(_workspace(n, k, dy.device)), ptr(dw),
                              ptr(db) if want_db else None, c_i(1 if accumulate else 0), stream_ptr(dy))
    check(lib, rc, "afl_tc_linear_dw")
    return dw, db


class TcLinear(torch.autograd.Function):
    """``F.linear`` (optionally + ReLU) on tcgen05 for 2-D CUDA inputs."""

    @staticmethod
    def forward(ctx, x, w, b, relu: bool = False):
        x2 = x.reshape(-1, x.shape[-1]).contiguous()
        y = linear_fwd(x2, w.contiguous(), b, relu=relu)
        ctx.save_for_backward(x2, w, y if relu else None)
        ctx.relu, ctx.shape = relu, x.shape
        return y.reshape(*x.shape[:-1], w.shape[0])

    @staticmethod
    def backward(ctx, gy):
        x2, w, y = ctx.saved_tensors
        g = gy.reshape(-1, gy.shape[-1]).contiguous()
        if ctx.relu:
            g = g * (y > 0)
        dx = linear_dx(g, w.contiguous()).reshape(ctx.shape)
        dw, db = linear_dw(g, x2)
        return dx, dw, db, None
