"""Binding of the tcgen05 Gram kernel (filled in with csrc/gram_tcgen05.cu)."""
from __future__ import annotations

import torch

from . import native


def available() -> bool:
    try:
        return hasattr(native.cuda_lib(), "afl_gram_bf16")
    except Exception:
        return False


def gram_tcgen05(a: torch.Tensor, precision: str = "bf16") -> torch.Tensor:
    raise NotImplementedError
