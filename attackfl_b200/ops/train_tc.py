"""Python binding of the tensor-core trainer for the ICU ``TransformerModel`` (``afl_tc_train_round``, ``csrc/train_tc.cu``).

One CTA trains one client for a whole FL round: every contraction is a ``tcgen05.mma`` (TF32 operands, fp32 accumulate in
TMEM), weights are streamed by TMA from the client's parameter row, activations needed by the backward live in a
per-client stash in global memory (L2-resident).  It draws the same Philox dropout masks as the FFMA interpreter
(``afl_train_round``), which stays the CPU twin, the oracle and the path for every other model.

Reference parity: ``train_ICU`` (``/root/reference/client.py:75-111``) on ``TransformerModel``
(``/root/reference/src/Model.py:184-191,225-246``).
"""
from __future__ import annotations

import ctypes
from typing import List, Optional, Tuple

import torch

from ..layout import FlatLayout
from .native import check, lib_for, ptr, stream_ptr

c_i, c_f, c_i64, c_u64 = ctypes.c_int, ctypes.c_float, ctypes.c_int64, ctypes.c_uint64

_BRANCHES = ("vitals", "labs")


def supported(model_name: str, layout: FlatLayout, batch: int, module=None) -> bool:
    """Default-width TransformerModel (d_model 64, 4 heads, ff <= 8, inputs 7 | 16) and batch <= 128."""
    if model_name != "TransformerModel" or not (2 <= batch <= 128):
        return False
    bn = layout.by_name
    try:
        d = bn["vitals_dense.weight"].shape[0]
        ff = bn["vitals_transformer.ffn.0.weight"].shape[0]
        kv, kl = bn["vitals_dense.weight"].shape[1], bn["labs_dense.weight"].shape[1]
        ok = (d == 64 and 1 <= ff <= 8 and kv <= 8 and kl <= 16 and bn["fc1.weight"].shape == (64, 128)
              and bn["fc2.weight"].shape == (32, 64) and bn["output.weight"].shape == (1, 32))
    except KeyError:
        return False
    heads = module.vitals_transformer.attention.num_heads if module is not None else 4
    return bool(ok and heads == 4)


def build_layout(layout: FlatLayout, dropout: bool) -> Tuple[List[int], List[int]]:
    """(layout words of ``TcLay``, parameter ranges [lo, hi) that receive gradients, merged and 4-aligned)."""
    off = layout.offset
    d = 64
    per = {k: [] for k in ("wd", "bd", "wv", "bv", "wo", "bo", "g1", "be1", "wf0", "bf0", "wf3", "bf3", "g2", "be2", "g3", "be3")}
    for b in _BRANCHES:
        t = f"{b}_transformer"
        per["wd"].append(off(f"{b}_dense.weight")); per["bd"].append(off(f"{b}_dense.bias"))
        per["wv"].append(off(f"{t}.attention.in_proj_weight") + 2 * d * d)
        per["bv"].append(off(f"{t}.attention.in_proj_bias") + 2 * d)
        per["wo"].append(off(f"{t}.attention.out_proj.weight")); per["bo"].append(off(f"{t}.attention.out_proj.bias"))
        per["g1"].append(off(f"{t}.attention_norm.weight")); per["be1"].append(off(f"{t}.attention_norm.bias"))
        per["wf0"].append(off(f"{t}.ffn.0.weight")); per["bf0"].append(off(f"{t}.ffn.0.bias"))
        per["wf3"].append(off(f"{t}.ffn.3.weight")); per["bf3"].append(off(f"{t}.ffn.3.bias"))
        per["g2"].append(off(f"{t}.ffn_norm.weight")); per["be2"].append(off(f"{t}.ffn_norm.bias"))
        per["g3"].append(off(f"{b}_bn.weight")); per["be3"].append(off(f"{b}_bn.bias"))
    words: List[int] = []
    for k in ("wd", "bd", "wv", "bv", "wo", "bo", "g1", "be1", "wf0", "bf0", "wf3", "bf3", "g2", "be2", "g3", "be3"):
        words += per[k]
    words += [off("fc1.weight"), off("fc1.bias"), off("fc2.weight"), off("fc2.bias"), off("output.weight"), off("output.bias")]
    bn = layout.by_name
    words += [bn["vitals_dense.weight"].shape[1], bn["labs_dense.weight"].shape[1], bn["vitals_transformer.ffn.0.weight"].shape[0]]
    # RNG stream ids in the order models/programs.py hands them out (vitals branch, labs branch, head)
    if dropout:
        words += [1, 5, 2, 6, 3, 7, 4, 8, 9]
    else:
        words += [0] * 9
    # ranges: every tensor except the Q/K thirds of in_proj (never read at T = 1: exactly zero gradient)
    spans = []
    for e in layout.entries:
        lo, hi = e.offset, e.offset + e.numel
        if e.name.endswith("attention.in_proj_weight"):
            lo += 2 * d * d
        elif e.name.endswith("attention.in_proj_bias"):
            lo += 2 * d
        spans.append((lo // 4 * 4, (hi + 3) // 4 * 4))
    spans.sort()
    merged: List[List[int]] = []
    for lo, hi in spans:
        if merged and lo - merged[-1][1] <= 64:        # bridge alignment padding (its gradient is zero)
            merged[-1][1] = max(merged[-1][1], hi)
        else:
            merged.append([lo, hi])
    ranges = [x for span in merged for x in span]
    return words, ranges


def stash_floats_per_client(device) -> int:
    lib = lib_for(torch.empty(0, device=device))
    return int(lib.afl_tc_stash_floats_per_client())


def last_cluster(device) -> int:
    """CTAs per client the last ``train_round_tc`` call of this process used (1, 2, 4 or 8)."""
    lib = lib_for(torch.empty(0, device=device))
    return int(lib.afl_tc_last_cluster())


def train_round_tc(layout: FlatLayout, w: torch.Tensor, m: torch.Tensor, v: Optional[torch.Tensor], gw: torch.Tensor,
                   stash: torch.Tensor, data: torch.Tensor, idx: torch.Tensor, n_data: torch.Tensor,
                   client_ids: torch.Tensor, batch: int, epochs: int, *, optimizer: str = "adam", lr: float = 0.004,
                   betas=(0.9, 0.999), eps: float = 1e-8, momentum: float = 0.0, clip_proper: bool = False,
                   clip: float = 0.0, seed: int = 0, dropout: bool = True, p1: float = 0.1, p3: float = 0.3,
                   cache: Optional[dict] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """One round of local training for ``w.shape[0]`` clients in place; same contract as ``ops.train.train_round``.

    ``gw``: ``[C, d_pad]`` gradient rows (scratch); ``stash``: ``[C, stash_floats_per_client]`` activation stash (scratch,
    zero-initialised once by the caller).
    """
    C, d_pad = w.shape
    dev = w.device
    lib = lib_for(w)
    assert dev.type == "cuda", "the tensor-core trainer needs a GPU"
    assert idx.dtype == torch.int32 and idx.dim() == 3 and idx.is_contiguous()
    assert gw.shape[0] >= C and gw.shape[1] == d_pad and stash.shape[0] >= C
    key = ("lay", id(layout), bool(dropout))
    if cache is not None and key in cache:
        words, ranges = cache[key]
    else:
        words, ranges = build_layout(layout, dropout)
        if cache is not None:
            cache[key] = (words, ranges)
    lay = (ctypes.c_int32 * len(words))(*words)
    rng = (ctypes.c_int32 * len(ranges))(*ranges)
    loss = torch.zeros(C, 2, dtype=torch.float32, device=dev)
    steps = torch.zeros(C, dtype=torch.int32, device=dev)
    rc = lib.afl_tc_train_round(
        lay, c_i(len(words)), ptr(w), ptr(m), ptr(v) if v is not None else ctypes.c_void_p(0), ptr(gw), ptr(stash),
        c_i64(d_pad), ptr(data), c_i(data.shape[1]), ptr(idx), c_i64(idx.stride(0)), c_i64(idx.stride(1)), ptr(n_data),
        ptr(client_ids), c_i(C), c_i(batch), c_i(epochs), c_i(0 if optimizer == "adam" else 1), c_f(lr), c_f(betas[0]),
        c_f(betas[1]), c_f(eps), c_f(momentum), c_i(1 if clip_proper else 0), c_f(clip), c_u64(seed & (2 ** 64 - 1)),
        c_i(1 if dropout else 0), c_f(p1 if dropout else 0.0), c_f(p3 if dropout else 0.0), rng, c_i(len(ranges) // 2),
        ptr(loss), ptr(steps), stream_ptr(w))
    check(lib, rc, "afl_tc_train_round")
    return loss, steps
