"""Gram matrix ``A A^T`` of a short-and-wide matrix ``A [R, D]`` (R <= 256 rows, D = model size).

One Gram serves Min-Max / Min-Sum / Opt-Fang (pairwise-distance search, ``/root/reference/src/Utils.py:
30-49,108-113,142-147,176-185``), Krum (``:326-342``), ShieldFL (``server.py:310-350``) and FLTrust
(``server.py:703-743``) — SURVEY §2.9.  On a GPU this is the split-K tcgen05/TMEM kernel K3
(``csrc/gram_tcgen05.cu``): rows are cast to bf16 (callers pass *centred* rows), staged by
TMA with 128-byte swizzle, multiplied by ``tcgen05.mma`` into a TMEM accumulator, and the per-CTA split-K
partials are reduced in fp32.  ``precision='bf16x3'`` uses the hi/lo split (3 MMAs) for near-fp32 results;
``'fp32'`` is the plain torch matmul (oracle).
"""
from __future__ import annotations

import torch


def gram_torch(a: torch.Tensor) -> torch.Tensor:
    a = a.to(torch.float32)
    return a @ a.t()


def gram(a: torch.Tensor, precision: str = "tf32") -> torch.Tensor:
    # the tcgen05 kernel holds the whole R x R tile in one TMEM accumulator: R <= 256 (k + 2 rows for the attacks, one row
    # per reporter for Krum / ShieldFL / FLTrust); larger problems use the library matmul in fp32
    if a.is_cuda and precision != "fp32" and a.shape[0] <= 256:
        from . import gram_native
        if gram_native.available():
            return gram_native.gram_tcgen05(a, precision)
    return gram_torch(a)
