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


def gram_tcgen05(a: torch.Tensor, precision: str = "bf16") -> torch.Tensor:
    """``A A^T`` (fp32 ``[R, R]``) of fp32 rows ``A [R, D]`` (R <= 256) with bf16 tensor-core operands."""
    assert a.is_cuda and a.dim() == 2 and a.dtype == torch.float32
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
