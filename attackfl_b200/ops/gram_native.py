"""Binding of the tcgen05 split-K Gram kernel K3 (``csrc/gram_tcgen05.cu``, ``afl_gram_bf16``)."""
from __future__ import annotations

import ctypes
from typing import Dict, Tuple

import torch

from . import native
from .native import check, ptr, stream_ptr

_ws: Dict[Tuple[int, int, int], Tuple[torch.Tensor, torch.Tensor]] = {}


def available() -> bool:
    try:
        return hasattr(native.cuda_lib(), "afl_gram_bf16")
    except Exception:
        return False


def workspace(R: int, D: int, device) -> Tuple[torch.Tensor, torch.Tensor]:
    """(bf16 operand buffer, fp32 split-K partials) for an ``R x D`` problem, cached per device/shape."""
    dev = torch.device(device)
    key = (dev.index or 0, R, D)
    if key not in _ws:
        lib = native.cuda_lib()
        n16, nf = ctypes.c_int64(0), ctypes.c_int64(0)
        check(lib, lib.afl_gram_workspace(ctypes.c_int(R), ctypes.c_int64(D), ctypes.byref(n16), ctypes.byref(nf)),
              "afl_gram_workspace")
        _ws[key] = (torch.empty(n16.value, dtype=torch.bfloat16, device=dev),
                    torch.empty(nf.value, dtype=torch.float32, device=dev))
        if len(_ws) > 16:
            _ws.pop(next(iter(_ws)))
    return _ws[key]


_ws32: Dict[Tuple[int, int, int], torch.Tensor] = {}


def gram_tf32(a: torch.Tensor) -> torch.Tensor:
    """``A A^T`` with the fp32 rows consumed directly as TF32 operands (one pass over A, no operand copy)."""
    R, D = a.shape
    if a.stride(1) != 1 or a.stride(0) % 4 != 0 or a.data_ptr() % 16 != 0:
        a = a.contiguous()
        if a.stride(0) % 4 != 0:
            pad = torch.zeros(R, (D + 3) // 4 * 4, dtype=torch.float32, device=a.device)
            pad[:, :D] = a
            a = pad
    lib = native.cuda_lib()
    key = (a.device.index or 0, R, D)
    if key not in _ws32:
        nf = ctypes.c_int64(0)
        check(lib, lib.afl_gram_tf32_workspace(ctypes.c_int(R), ctypes.c_int64(D), ctypes.byref(nf)), "afl_gram_tf32_workspace")
        _ws32[key] = torch.empty(nf.value, dtype=torch.float32, device=a.device)
        if len(_ws32) > 16:
            _ws32.pop(next(iter(_ws32)))
    G = torch.empty(R, R, dtype=torch.float32, device=a.device)
    rc = lib.afl_gram_tf32(ctypes.c_void_p(a.data_ptr()), ctypes.c_int(R), ctypes.c_int64(D), ctypes.c_int64(a.stride(0)),
                           ptr(_ws32[key]), ptr(G), stream_ptr(a))
    check(lib, rc, "afl_gram_tf32")
    return G


def gram_tcgen05(a: torch.Tensor, precision: str = "tf32") -> torch.Tensor:
    """``A A^T`` (fp32 ``[R, R]``) of fp32 rows ``A [R, D]`` (R <= 256) on tcgen05 tensor cores.

    ``tf32`` (default): fp32 rows read directly as TF32 operands; ``bf16``: cast pass + bf16 operands;
    ``bf16x3``: hi/lo split for near-fp32 results."""
    assert a.is_cuda and a.dim() == 2 and a.dtype == torch.float32
    if precision == "tf32":
        return gram_tf32(a)
    if precision == "bf16x3":
        # hi/lo split: A = H + L with H, L exactly representable in bf16
        return _gram_split(a, a.to(torch.bfloat16).to(torch.float32))
    a = a if a.stride(1) == 1 else a.contiguous()
    R, D = a.shape
    lib = native.cuda_lib()
    a16, part = workspace(R, D, a.device)
    G = torch.empty(R, R, dtype=torch.float32, device=a.device)
    rc = lib.afl_gram_bf16(ptr(a) if a.is_contiguous() else ctypes.c_void_p(a.data_ptr()), ctypes.c_int(R),
                           ctypes.c_int64(D), ctypes.c_int64(a.stride(0)), ptr(a16), ptr(part), ptr(G), stream_ptr(a))
    check(lib, rc, "afl_gram_bf16")
    return G


def _gram_split(a: torch.Tensor, h: torch.Tensor) -> torch.Tensor:
    """3-pass exactness mode: rows [H; L] stacked (2R <= 256) give HH^T, HL^T, LL^T blocks in one Gram."""
    R = a.shape[0]
    if 2 * R > 256:
        return (a @ a.t())
    g = gram_tcgen05(torch.cat([h, a - h], dim=0).contiguous(), "bf16")
    return g[:R, :R] + g[:R, R:] + g[R:, :R] + g[R:, R:]
