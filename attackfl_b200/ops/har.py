"""Native HAR trainer: one optimiser step of ``TransformerClassifier`` (T = 561, d = 64, 4 heads, 2 layers) run
entirely by this repo's sm_100a kernels on a row of the flat parameter arena.

* linear layers (in_proj / out_proj / linear1 / linear2, forward + both gradients): ``tcgen05`` TF32 GEMMs with fused
  bias / ReLU / dropout / mask epilogues (``csrc/gemm_tcgen05.cu``);
* Conv1d embedding + positional encoding, flash-style attention with attention-weight dropout, residual + dropout +
  LayerNorm, mean pooling, classifier head + cross-entropy — forward and backward (``csrc/har_kernels.cu``);
* gradients land directly in a flat vector with the arena layout, so the step ends with the fused flat Adam / SGD kernel.

Reference counterpart: ``train_HAR`` (``/root/reference/client.py:114-131``) on ``TransformerClassifier``
(``/root/reference/src/Model.py:435-458``).  No host synchronisation happens inside a step; the loss sum, the NaN
count and the number of correct predictions accumulate in a 3-float device buffer.
"""
from __future__ import annotations

import ctypes
import os
from typing import Dict, Optional

import torch

from ..layout import FlatLayout
from . import tc_linear as tc
from .native import check, cuda_lib, ptr

c_i, c_f, c_i64, c_u64, c_u32 = ctypes.c_int, ctypes.c_float, ctypes.c_int64, ctypes.c_uint64, ctypes.c_uint32

D_MODEL, N_HEADS, FF, N_LAYERS = 64, 4, 256, 2
P_ENC, P_HEAD = 0.1, 0.3
# dropout sites (Philox stream ids); attention uses site + head
SITE_ATTN, SITE_SA, SITE_FF, SITE_OUT, SITE_HEAD = 0, 8, 9, 10, 100


def supports(layout: FlatLayout) -> bool:
    names = {e.name for e in layout.entries}
    return "transformer.layers.1.linear2.weight" in names and "classifier.3.weight" in names and "conv.weight" in names


class HarNativeTrainer:
    def __init__(self, layout: FlatLayout, device, max_batch: int = 128, seq_len: int = 561, n_classes: int = 6):
        self.layout, self.device = layout, torch.device(device)
        self.lib = cuda_lib()
        self.T, self.n_cls, self.max_batch = seq_len, n_classes, max_batch
        # attention forward: tcgen05 kernel when K/V of a head pair fit in shared memory, else the FFMA kernel
        mode = os.environ.get("ATTACKFL_HAR_ATTN", "tc")            # tc | tc-fwd (FFMA backward) | ffma
        self.attn_tc = bool(self.lib.afl_har_attn_tc_supported(c_i(seq_len))) and mode != "ffma"
        self.attn_bwd_tc = mode == "tc"
        self.fuse_ln = os.environ.get("ATTACKFL_HAR_FUSE_LN", "1") != "0"      # LayerNorm in the GEMM epilogue
        m = max_batch * seq_len
        z = lambda *s: torch.empty(*s, dtype=torch.float32, device=self.device)          # noqa: E731
        self.h0 = z(m, D_MODEL)
        self.Tp = int(self.lib.afl_har_lse_stride(c_i(seq_len)))
        self.mask_words = int(self.lib.afl_har_attn_mask_words(c_i(seq_len)))
        zz = lambda n, dt=torch.float32: torch.zeros(n, dtype=dt, device=self.device)       # noqa: E731
        self.delta = zz(max_batch * N_HEADS * self.Tp)
        self.layers = [{"qkv": z(m, 3 * D_MODEL), "att": z(m, D_MODEL), "lse": zz(max_batch * N_HEADS * self.Tp),
                        "mask": zz(max_batch * N_HEADS * self.mask_words * self.Tp, torch.int32) if self.attn_tc else None,
                        "sa": z(m, D_MODEL), "h1": z(m, D_MODEL), "xhat1": z(m, D_MODEL), "rstd1": z(m),
                        "f1": z(m, FF), "f2": z(m, D_MODEL), "h2": z(m, D_MODEL), "xhat2": z(m, D_MODEL), "rstd2": z(m)}
                       for _ in range(N_LAYERS)]
        self.dA, self.dB, self.dC, self.dD = (z(m, D_MODEL) for _ in range(4))
        self.d192, self.d256 = z(m, 3 * D_MODEL), z(m, FF)
        self.pooled, self.dpooled = z(max_batch, D_MODEL), z(max_batch, D_MODEL)
        self.logits = z(max_batch, n_classes)
        self.out3 = torch.zeros(3, dtype=torch.float32, device=self.device)

    # ------------------------------------------------------------------------------------------------ helpers
    def _st(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)

    def _ck(self, rc, what):
        check(self.lib, rc, what)

    def _views(self, flat: torch.Tensor) -> Dict[str, torch.Tensor]:
        return self.layout.views(flat)

    def forward(self, w: Dict[str, torch.Tensor], x: torch.Tensor, seed: int, client: int, step: int, train: bool):
        """Runs the encoder on ``x [B, 561]`` and leaves the activations in the trainer's buffers."""
        lib, st, T = self.lib, self._st(), self.T
        B = x.shape[0]
        M = B * T
        p = P_ENC if train else 0.0
        self._ck(lib.afl_har_embed_fwd(ptr(x), ptr(w["conv.weight"]), ptr(w["conv.bias"]), ptr(w["pe.pe"]), c_i(B), c_i(T),
                                       ptr(self.h0), st), "afl_har_embed_fwd")
        h = self.h0
        for l, a in enumerate(self.layers):
            pre = f"transformer.layers.{l}."
            rng = (client, step)
            tc.linear_fwd(h[:M], w[pre + "self_attn.in_proj_weight"], w[pre + "self_attn.in_proj_bias"], out=a["qkv"][:M])
            if self.attn_tc:
                self._ck(lib.afl_har_attn_fwd_tc(ptr(a["qkv"]), c_i(B), c_i(T), c_f(p), c_u64(seed), c_u32(client),
                                                 c_u32(step), c_u32(16 * l + SITE_ATTN), ptr(a["att"]), ptr(a["lse"]),
                                                 ptr(a["mask"]) if train else None, st), "afl_har_attn_fwd_tc")
            else:
                self._ck(lib.afl_har_attn_fwd(ptr(a["qkv"]), c_i(B), c_i(T), c_f(p), c_u64(seed), c_u32(client), c_u32(step),
                                              c_u32(16 * l + SITE_ATTN), ptr(a["att"]), ptr(a["lse"]), st), "afl_har_attn_fwd")
            if self.fuse_ln:        # out_proj with residual + dropout + LayerNorm in its epilogue
                tc.linear_ln_fwd(a["att"][:M], w[pre + "self_attn.out_proj.weight"], w[pre + "self_attn.out_proj.bias"],
                                 h[:M], w[pre + "norm1.weight"], w[pre + "norm1.bias"], a["h1"][:M], a["xhat1"][:M],
                                 a["rstd1"][:M], drop_p=p, seed=seed, rng=(client, step, 16 * l + SITE_SA))
            else:
                tc.linear_fwd(a["att"][:M], w[pre + "self_attn.out_proj.weight"], w[pre + "self_attn.out_proj.bias"],
                              out=a["sa"][:M])
                self._ck(lib.afl_har_add_ln_fwd(ptr(h), ptr(a["sa"]), c_f(p), c_u64(seed), c_u32(client), c_u32(step),
                                                c_u32(16 * l + SITE_SA), ptr(w[pre + "norm1.weight"]),
                                                ptr(w[pre + "norm1.bias"]), c_i64(M), ptr(a["h1"]), ptr(a["xhat1"]),
                                                ptr(a["rstd1"]), st), "afl_har_add_ln_fwd")
            tc.linear_fwd(a["h1"][:M], w[pre + "linear1.weight"], w[pre + "linear1.bias"], relu=True, drop_p=p, seed=seed,
                          rng=(rng[0], rng[1], 16 * l + SITE_FF), out=a["f1"][:M])
            if self.fuse_ln:
                tc.linear_ln_fwd(a["f1"][:M], w[pre + "linear2.weight"], w[pre + "linear2.bias"], a["h1"][:M],
                                 w[pre + "norm2.weight"], w[pre + "norm2.bias"], a["h2"][:M], a["xhat2"][:M], a["rstd2"][:M],
                                 drop_p=p, seed=seed, rng=(client, step, 16 * l + SITE_OUT))
            else:
                tc.linear_fwd(a["f1"][:M], w[pre + "linear2.weight"], w[pre + "linear2.bias"], out=a["f2"][:M])
                self._ck(lib.afl_har_add_ln_fwd(ptr(a["h1"]), ptr(a["f2"]), c_f(p), c_u64(seed), c_u32(client), c_u32(step),
                                                c_u32(16 * l + SITE_OUT), ptr(w[pre + "norm2.weight"]),
                                                ptr(w[pre + "norm2.bias"]), c_i64(M), ptr(a["h2"]), ptr(a["xhat2"]),
                                                ptr(a["rstd2"]), st), "afl_har_add_ln_fwd")
            h = a["h2"]
        self._ck(lib.afl_har_pool_fwd(ptr(h), c_i(B), c_i(T), ptr(self.pooled), st), "afl_har_pool_fwd")

    def _head(self, w, g, y: torch.Tensor, B: int, seed: int, client: int, step: int, train: bool):
        gp = (lambda k: ptr(g[k])) if g is not None else (lambda k: None)
        self._ck(self.lib.afl_har_head(ptr(self.pooled), ptr(w["classifier.0.weight"]), ptr(w["classifier.0.bias"]),
                                       ptr(w["classifier.3.weight"]), ptr(w["classifier.3.bias"]), ptr(y), c_i(B),
                                       c_i(self.n_cls), c_f(P_HEAD if train else 0.0), c_u64(seed), c_u32(client),
                                       c_u32(step), c_u32(SITE_HEAD), c_i(1 if g is not None else 0), ptr(self.logits),
                                       ptr(self.out3), ptr(self.dpooled) if g is not None else None,
                                       gp("classifier.0.weight"), gp("classifier.0.bias"), gp("classifier.3.weight"),
                                       gp("classifier.3.bias"), self._st()), "afl_har_head")

    # ------------------------------------------------------------------------------------------------ public
    def logits_of(self, row: torch.Tensor, x: torch.Tensor, y: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Inference logits ``[B, n_cls]`` (no dropout) for ``x [B, 561]``."""
        B = x.shape[0]
        w = self._views(row)
        self.forward(w, x.contiguous(), 0, 0, 0, train=False)
        yy = y if y is not None else torch.zeros(B, dtype=torch.int64, device=self.device)
        self._head(w, None, yy, B, 0, 0, 0, train=False)
        return self.logits[:B]

    def grad_step(self, row: torch.Tensor, grad: torch.Tensor, x: torch.Tensor, y: torch.Tensor, seed: int = 0,
                  client: int = 0, step: int = 0, dropout: bool = True) -> torch.Tensor:
        """Forward + backward of one batch: writes the mean-loss gradient into ``grad`` (arena layout) and adds
        ``{loss sum, non-finite count, correct}`` to ``self.out3`` (returned)."""
        lib, st, T = self.lib, self._st(), self.T
        B = x.shape[0]
        M = B * T
        assert B <= self.max_batch and x.shape[1] == T
        x = x.contiguous()
        w, g = self._views(row), self._views(grad)
        grad.zero_()
        p = P_ENC if dropout else 0.0
        self.forward(w, x, seed, client, step, train=dropout)
        self._head(w, g, y, B, seed, client, step, train=dropout)
        dA, dB, dC, dD = self.dA, self.dB, self.dC, self.dD
        dh = None
        for l in reversed(range(N_LAYERS)):
            a = self.layers[l]
            pre = f"transformer.layers.{l}."
            h_in = self.h0 if l == 0 else self.layers[l - 1]["h2"]
            # LayerNorm 2 (+ the mean-pool backward for the last layer)
            self._ck(lib.afl_har_ln_bwd(ptr(dh) if dh is not None else None, ptr(self.dpooled) if dh is None else None,
                                        c_i(T), c_f(1.0 / T), ptr(a["xhat2"]), ptr(a["rstd2"]), ptr(w[pre + "norm2.weight"]),
                                        c_f(p), c_u64(seed), c_u32(client), c_u32(step), c_u32(16 * l + SITE_OUT), c_i64(M),
                                        ptr(dA), ptr(dB), ptr(g[pre + "norm2.weight"]), ptr(g[pre + "norm2.bias"]), st),
                     "afl_har_ln_bwd")
            # feed-forward block: dB = d f2, dA = d h1 (residual path)
            tc.linear_dw(dB[:M], a["f1"][:M], dw=g[pre + "linear2.weight"], db=g[pre + "linear2.bias"])
            tc.linear_dx(dB[:M], w[pre + "linear2.weight"], mask_src=a["f1"][:M], mask_scale=1.0 / (1.0 - p),
                         out=self.d256[:M])
            tc.linear_dw(self.d256[:M], a["h1"][:M], dw=g[pre + "linear1.weight"], db=g[pre + "linear1.bias"])
            tc.linear_dx(self.d256[:M], w[pre + "linear1.weight"], out=dA[:M], accumulate=True)
            # LayerNorm 1: dC = d h_in (residual path), dD = d sa
            self._ck(lib.afl_har_ln_bwd(ptr(dA), None, c_i(T), c_f(1.0), ptr(a["xhat1"]), ptr(a["rstd1"]),
                                        ptr(w[pre + "norm1.weight"]), c_f(p), c_u64(seed), c_u32(client), c_u32(step),
                                        c_u32(16 * l + SITE_SA), c_i64(M), ptr(dC), ptr(dD), ptr(g[pre + "norm1.weight"]),
                                        ptr(g[pre + "norm1.bias"]), st), "afl_har_ln_bwd")
            tc.linear_dw(dD[:M], a["att"][:M], dw=g[pre + "self_attn.out_proj.weight"], db=g[pre + "self_attn.out_proj.bias"])
            tc.linear_dx(dD[:M], w[pre + "self_attn.out_proj.weight"], out=dB[:M])
            if self.attn_tc and self.attn_bwd_tc:
                self._ck(lib.afl_har_attn_bwd_tc(ptr(a["qkv"]), ptr(a["att"]), ptr(dB), ptr(a["lse"]), ptr(a["mask"]), c_i(B),
                                                 c_i(T), c_f(p), ptr(self.delta), ptr(self.d192), st), "afl_har_attn_bwd_tc")
            else:
                self._ck(lib.afl_har_attn_bwd(ptr(a["qkv"]), ptr(a["att"]), ptr(dB), ptr(a["lse"]), c_i(B), c_i(T), c_f(p),
                                              c_u64(seed), c_u32(client), c_u32(step), c_u32(16 * l + SITE_ATTN),
                                              ptr(self.d192), st), "afl_har_attn_bwd")
            tc.linear_dw(self.d192[:M], h_in[:M], dw=g[pre + "self_attn.in_proj_weight"],
                         db=g[pre + "self_attn.in_proj_bias"])
            tc.linear_dx(self.d192[:M], w[pre + "self_attn.in_proj_weight"], out=dC[:M], accumulate=True)
            dh = dC      # consumed by the next iteration's LayerNorm-2 backward before dC is written again
        self._ck(lib.afl_har_embed_bwd(ptr(dh), ptr(x), c_i(B), c_i(T), ptr(g["conv.weight"]), ptr(g["conv.bias"]), st),
                 "afl_har_embed_bwd")
        return self.out3
