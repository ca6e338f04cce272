"""Op-program builders: each ICU model as the instruction list of the persistent training kernel.

The kernel (``csrc/train_core.h``) interprets a short program of fused ops on tiles of ``S`` samples
whose activations live in one shared-memory pool.  This module (i) emits the ops for ``CNNModel``,
``RNNModel`` and ``TransformerModel`` with parameter offsets taken from the ``FlatLayout`` — the
closed forms of SURVEY §2.7: GRU at T=1 with h0=0, attention over one token = out_proj(V x) with
head-level dropout, AdaptiveAvgPool1d(4) windows — and (ii) plans the pool: activations are
allocated once, gradient buffers are packed by liveness (a gradient buffer lives from the backward
of its last consumer to the backward of its first producer).

Reference semantics implemented: ``/root/reference/src/Model.py:56-88`` (CNN forward), ``:120-163``
(RNN forward), ``:184-191,225-246`` (Transformer block / model forward), loss ``client.py:77,95``.
"""
from __future__ import annotations

import struct
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Tuple

import numpy as np
import torch

from ..layout import FlatLayout
from .icu import LABS_DIM, VITALS_DIM

OP_WORDS = 32
LN_PARTS = 8
K_LOAD, K_LINEAR, K_LN, K_GRU, K_CONV, K_POOL, K_DROPOUT, K_BCE = 1, 2, 3, 4, 5, 6, 7, 8
ACT_NONE, ACT_RELU, ACT_GELU = 0, 1, 2
F_ACC0, F_ACC1, F_NEED_GIN0, F_MASK_INPUT = 1, 2, 4, 8

_FIELDS = ["kind", "act", "in0", "in0_ld", "in1", "in1_ld", "out", "out_ld", "aux", "aux2", "w", "b", "w2",
           "N", "K", "L", "p", "rng", "g_out", "g_in0", "g_in1", "flags", "in_cs", "out_cs", "grp_start", "grp_size"]


def _align(x: int, a: int = 4) -> int:
    return (x + a - 1) // a * a


@dataclass
class Buf:
    """A view ``[S, width]`` (row stride ``ld``) into a root buffer of the pool."""
    root: int            # root id
    col: int             # offset inside a root row
    ld: int
    width: int


@dataclass
class _Root:
    size: int            # floats (S * ld, or explicit)
    off: int = -1        # activation offset
    goff: int = -1       # gradient offset
    producers: List[int] = field(default_factory=list)
    consumers: List[int] = field(default_factory=list)
    needs_grad: bool = True


@dataclass
class Program:
    ops: torch.Tensor            # int32 [n_ops, 32]
    n_ops: int
    pool_floats: int
    S: int
    rec: int
    model: str
    train: bool
    act_floats: int = 0
    grad_floats: int = 0
    chain: bool = False          # planned for chain mode (no gradient-buffer reuse)

    def to(self, device) -> "Program":
        return Program(self.ops.to(device), self.n_ops, self.pool_floats, self.S, self.rec, self.model,
                       self.train, self.act_floats, self.grad_floats, self.chain)

    def smem_bytes(self, threads: int = 512) -> int:
        table = 4 * OP_WORDS * self.n_ops * (1 + (self.S if self.chain else 0))
        return 4 * self.pool_floats + table + 4 * (threads + 4)


class Builder:
    def __init__(self, layout: FlatLayout, S: int, train: bool = True, reuse_grads: bool = True):
        self.layout, self.S, self.train, self.reuse_grads = layout, S, train, reuse_grads
        self.roots: List[_Root] = []
        self.ops: List[Dict] = []
        self._sink: List[Dict] = self.ops
        self._rng = 0

    # -- buffers
    def new(self, width: int, ld: Optional[int] = None, needs_grad: bool = True, rows: Optional[int] = None) -> Buf:
        ld = ld if ld is not None else width
        rows = self.S if rows is None else rows
        self.roots.append(_Root(size=_align(rows * ld), needs_grad=needs_grad))
        return Buf(len(self.roots) - 1, 0, ld, width)

    def view(self, b: Buf, col: int, width: int) -> Buf:
        return Buf(b.root, b.col + col, b.ld, width)

    def _use(self, op: Dict, produced: List[Optional[Buf]], consumed: List[Optional[Buf]]):
        op["_produced"] = [b for b in produced if b is not None]
        op["_consumed"] = [b for b in consumed if b is not None]

    def _emit(self, **kw) -> Dict:
        self._sink.append(kw)
        return kw

    # -- concurrent groups: independent op sequences (e.g. the two input branches) are emitted into separate
    #    lists and zipped; ops at the same position form a group the kernel runs in one barrier interval.
    def branch(self, fn) -> List[Dict]:
        keep, self._sink = self._sink, []
        try:
            fn()
            return self._sink
        finally:
            self._sink = keep

    def zip_groups(self, *lists: List[Dict]) -> None:
        n = len(lists[0])
        assert all(len(l) == n for l in lists), "branches must have the same op sequence"
        for j in range(n):
            members = [l[j] for l in lists]
            assert len({m["kind"] for m in members}) == 1, "grouped ops must have the same kind"
            touched = set()
            for m in members:                      # gradient targets must be disjoint inside a group (no races)
                for b in m["_consumed"]:
                    if self.roots[b.root].needs_grad:
                        assert b.root not in touched, "ops of one group write the same gradient buffer"
                        touched.add(b.root)
            for m in members:
                m["_group"] = members
                self._sink.append(m)

    def _next_rng(self) -> int:
        self._rng += 1
        return self._rng

    def off(self, key: str) -> int:
        return self.layout.offset(key)

    # -- ops
    def load(self, mask_input: bool) -> Tuple[Buf, Buf, Buf]:
        v = self.new(VITALS_DIM, ld=_align(VITALS_DIM), needs_grad=False)
        l = self.new(LABS_DIM, needs_grad=False)
        y = self.new(1, needs_grad=False)
        i = self._emit(kind=K_LOAD, out=v, aux=l, in1=y, N=VITALS_DIM, K=LABS_DIM,
                       flags=F_MASK_INPUT if mask_input else 0)
        self._use(i, [v, l, y], [])
        return v, l, y

    def linear(self, x: Buf, wkey: str, bkey: str, N: int, K: int, act: int = ACT_NONE, p: float = 0.0,
               w_row0: int = 0, out: Optional[Buf] = None) -> Buf:
        out = out if out is not None else self.new(N)
        aux = self.new(N, needs_grad=False) if act == ACT_GELU else None
        i = self._emit(kind=K_LINEAR, act=act, in0=x, out=out, aux=aux, w=self.off(wkey) + w_row0 * K,
                       b=self.off(bkey) + w_row0, N=N, K=K, p=p, rng=self._next_rng() if p > 0 else 0)
        self._use(i, [out], [x])
        return out

    def layernorm(self, x: Buf, wkey: str, bkey: str, N: int, res: Optional[Buf] = None, p: float = 0.0,
                  out: Optional[Buf] = None) -> Buf:
        out = out if out is not None else self.new(N)
        xhat = self.new(N, needs_grad=False)
        part = self.new(1, ld=1, needs_grad=False, rows=self.S * (2 * LN_PARTS + 1))
        i = self._emit(kind=K_LN, in0=x, in1=res, out=out, aux=xhat, aux2=part, w=self.off(wkey), b=self.off(bkey),
                       N=N, p=p, rng=self._next_rng() if p > 0 else 0)
        self._use(i, [out], [x, res])
        return out

    def gru(self, gates: Buf, bhh_key: str, H: int, out: Buf) -> Buf:
        i = self._emit(kind=K_GRU, in0=gates, out=out, w2=self.off(bhh_key), N=H)
        self._use(i, [out], [gates])
        return out

    def conv(self, x: Buf, wkey: str, bkey: str, Cin: int, Cout: int, L: int, in_cs: int) -> Buf:
        Lp = L | 1                                  # odd channel stride: conflict-free column walks
        out = self.new(Cout * Lp)
        i = self._emit(kind=K_CONV, in0=x, out=out, w=self.off(wkey), b=self.off(bkey), N=Cout, K=Cin, L=L,
                       in_cs=in_cs, out_cs=Lp)
        self._use(i, [out], [x])
        return out

    def pool(self, x: Buf, C: int, L: int, in_cs: int, p: float, out: Buf) -> Buf:
        i = self._emit(kind=K_POOL, in0=x, out=out, N=C, L=L, in_cs=in_cs, p=p,
                       rng=self._next_rng() if p > 0 else 0)
        self._use(i, [out], [x])
        return out

    def dropout(self, x: Buf, N: int, p: float, group: int = 1, out: Optional[Buf] = None) -> Buf:
        out = out if out is not None else self.new(N)
        i = self._emit(kind=K_DROPOUT, in0=x, out=out, N=N, L=group, p=p, rng=self._next_rng())
        self._use(i, [out], [x])
        return out

    def bce(self, logit: Buf, labels: Buf) -> None:
        scratch = self.new(1, needs_grad=False)
        i = self._emit(kind=K_BCE, in0=logit, in1=labels, out=scratch)
        self._use(i, [scratch], [logit])

    # -- planning + packing
    def finish(self, model: str) -> Program:
        for r in self.roots:
            r.producers, r.consumers = [], []
        for i, op in enumerate(self.ops):
            op["_idx"] = i
            for b in op.get("_produced", []):
                self.roots[b.root].producers.append(i)
            for b in op.get("_consumed", []):
                self.roots[b.root].consumers.append(i)
        off = 0
        for r in self.roots:
            r.off = off
            off += r.size
        act_floats = off
        grad_floats = 0
        if self.train:
            # interval of a gradient buffer in op-index space: [first producer, last consumer]
            live = []
            for rid, r in enumerate(self.roots):
                if not r.needs_grad or not r.consumers:
                    continue
                # ops of a concurrent group share one time step: widen to the group's first / last member
                hi_op, lo_op = self.ops[max(r.consumers)], self.ops[min(r.producers) if r.producers else 0]
                hi_t = hi_op["_group"][-1]["_idx"] if hi_op.get("_group") else hi_op["_idx"]
                lo_t = lo_op["_group"][0]["_idx"] if lo_op.get("_group") else lo_op["_idx"]
                live.append((hi_t, lo_t, rid))
            if not self.reuse_grads:
                # chain mode defers the parameter gradients to the end of the tile, so every gradient buffer must
                # survive until then: give all of them the whole program as lifetime
                live = [(len(self.ops), 0, rid) for _, _, rid in live]
            live.sort(reverse=True)
            placed: List[Tuple[int, int, int, int]] = []      # (goff, size, lo, hi)
            for hi, lo, rid in live:
                size = self.roots[rid].size
                cand = 0
                for g0, gs, l0, h0 in sorted(placed):
                    overlap_time = not (hi < l0 or h0 < lo)
                    if overlap_time and cand + size > g0 and cand < g0 + gs:
                        cand = g0 + gs
                self.roots[rid].goff = cand
                placed.append((cand, size, lo, hi))
                grad_floats = max(grad_floats, cand + size)
        pool_floats = _align(act_floats + grad_floats, 4)

        def a_off(b: Optional[Buf]) -> int:
            return -1 if b is None else self.roots[b.root].off + b.col

        def g_off(b: Optional[Buf]) -> int:
            if b is None or not self.train:
                return 0
            r = self.roots[b.root]
            return 0 if r.goff < 0 else act_floats + r.goff + b.col

        words = np.zeros((len(self.ops), OP_WORDS), dtype=np.int32)
        for i, op in enumerate(self.ops):
            rec = {k: 0 for k in _FIELDS}
            rec.update(kind=op["kind"], act=op.get("act", 0), w=op.get("w", -1), b=op.get("b", -1),
                       w2=op.get("w2", -1), N=op.get("N", 0), K=op.get("K", 0), L=op.get("L", 0),
                       rng=op.get("rng", 0), in_cs=op.get("in_cs", 0), out_cs=op.get("out_cs", 0))
            in0, in1, out = op.get("in0"), op.get("in1"), op.get("out")
            rec["in0"], rec["in0_ld"] = a_off(in0), (in0.ld if in0 else 0)
            rec["in1"], rec["in1_ld"] = a_off(in1), (in1.ld if in1 else 0)
            rec["out"], rec["out_ld"] = a_off(out), (out.ld if out else 0)
            aux, aux2 = op.get("aux"), op.get("aux2")
            rec["aux"] = a_off(aux) if isinstance(aux, Buf) else -1
            if op["kind"] == K_LOAD:
                rec["aux2"] = aux.ld
            else:
                rec["aux2"] = a_off(aux2) if isinstance(aux2, Buf) else -1
            flags = op.get("flags", 0)
            if self.train:
                rec["g_out"] = g_off(out)
                if in0 is not None and self.roots[in0.root].needs_grad:
                    flags |= F_NEED_GIN0
                    rec["g_in0"] = g_off(in0)
                    if i != max(self.roots[in0.root].consumers):
                        flags |= F_ACC0
                if in1 is not None and op["kind"] == K_LN:
                    rec["g_in1"] = g_off(in1)
                    if i != max(self.roots[in1.root].consumers):
                        flags |= F_ACC1
                if op["kind"] == K_BCE:
                    rec["g_in0"] = g_off(in0)
            rec["flags"] = flags
            grp = op.get("_group")
            if grp:
                rec["grp_start"] = grp[0]["_idx"]
                rec["grp_size"] = len(grp) if op is grp[0] else 0
            else:
                rec["grp_start"], rec["grp_size"] = i, 1
            for j, k in enumerate(_FIELDS):
                if k == "p":
                    words[i, j] = struct.unpack("<i", struct.pack("<f", float(op.get("p", 0.0))))[0]
                else:
                    words[i, j] = int(rec[k])
        return Program(torch.from_numpy(words).contiguous(), len(self.ops), pool_floats, self.S,
                       VITALS_DIM + LABS_DIM + 1, model, self.train, act_floats, grad_floats)


# ---------------------------------------------------------------------------------------- models
def _transformer(bd: Builder, d: int, ff: int, heads: int, dropout: bool):
    p1, p3 = (0.1, 0.3) if dropout else (0.0, 0.0)
    v, l, y = bd.load(mask_input=False)
    cat = bd.new(2 * d)

    def branch(bi, name, x, k_in):
        t = f"{name}_transformer"
        h0 = bd.linear(x, f"{name}_dense.weight", f"{name}_dense.bias", d, k_in, ACT_GELU)
        # one token => softmax == 1 => attention = out_proj(dropout_head(V h0)); Q/K never touched
        val = bd.linear(h0, f"{t}.attention.in_proj_weight", f"{t}.attention.in_proj_bias", d, d, w_row0=2 * d)
        if p1 > 0:
            val = bd.dropout(val, d, p1, group=d // heads)
        att = bd.linear(val, f"{t}.attention.out_proj.weight", f"{t}.attention.out_proj.bias", d, d)
        y1 = bd.layernorm(h0, f"{t}.attention_norm.weight", f"{t}.attention_norm.bias", d, res=att, p=p1)
        f1 = bd.linear(y1, f"{t}.ffn.0.weight", f"{t}.ffn.0.bias", ff, d, ACT_GELU, p=p1)
        f2 = bd.linear(f1, f"{t}.ffn.3.weight", f"{t}.ffn.3.bias", d, ff)
        y2 = bd.layernorm(y1, f"{t}.ffn_norm.weight", f"{t}.ffn_norm.bias", d, res=f2, p=p1)
        bd.layernorm(y2, f"{name}_bn.weight", f"{name}_bn.bias", d, out=bd.view(cat, bi * d, d))

    bd.zip_groups(bd.branch(lambda: branch(0, "vitals", v, VITALS_DIM)),
                  bd.branch(lambda: branch(1, "labs", l, LABS_DIM)))
    c1 = bd.linear(cat, "fc1.weight", "fc1.bias", d, 2 * d, ACT_GELU, p=p3)
    c2 = bd.linear(c1, "fc2.weight", "fc2.bias", d // 2, d, ACT_GELU)
    logit = bd.linear(c2, "output.weight", "output.bias", 1, d // 2)
    bd.bce(logit, y)


def _rnn(bd: Builder, H: int, dropout: bool):
    p3 = 0.3 if dropout else 0.0
    v, l, y = bd.load(mask_input=True)
    cat = bd.new(4 * H)

    def branch(bi, name, x, k_in):
        cur, width = x, k_in
        for layer in (1, 2, 3):
            out = bd.new(2 * H)
            for di, suffix in enumerate(("", "_reverse")):
                g = f"{name}_gru{layer}"
                gates = bd.linear(cur, f"{g}.weight_ih_l0{suffix}", f"{g}.bias_ih_l0{suffix}", 3 * H, width)
                bd.gru(gates, f"{g}.bias_hh_l0{suffix}", H, bd.view(out, di * H, H))
            cur, width = out, 2 * H
        tgt = bd.view(cat, bi * 2 * H, 2 * H)
        if p3 > 0:
            ln = bd.layernorm(cur, f"{name}_ln.weight", f"{name}_ln.bias", 2 * H)
            bd.dropout(ln, 2 * H, p3, out=tgt)
        else:
            bd.layernorm(cur, f"{name}_ln.weight", f"{name}_ln.bias", 2 * H, out=tgt)

    bd.zip_groups(bd.branch(lambda: branch(0, "vitals", v, VITALS_DIM)),
                  bd.branch(lambda: branch(1, "labs", l, LABS_DIM)))
    c1 = bd.linear(cat, "fc1.weight", "fc1.bias", H, 4 * H, ACT_RELU)
    c2 = bd.linear(c1, "fc2.weight", "fc2.bias", H // 2, H, ACT_RELU)
    logit = bd.linear(c2, "output.weight", "output.bias", 1, H // 2)
    bd.bce(logit, y)


def _cnn(bd: Builder, dropout: bool):
    p3 = 0.3 if dropout else 0.0
    ch = (1, 32, 64, 128)
    v, l, y = bd.load(mask_input=False)
    cat = bd.new(2 * ch[3] * 4)

    def branch(bi, name, x, L):
        cur, cs = x, L
        for i in range(3):
            cur = bd.conv(cur, f"{name}_conv{i + 1}.weight", f"{name}_conv{i + 1}.bias", ch[i], ch[i + 1], L, cs)
            cs = L | 1
        bd.pool(cur, ch[3], L, cs, p3, bd.view(cat, bi * ch[3] * 4, ch[3] * 4))

    bd.zip_groups(bd.branch(lambda: branch(0, "vitals", v, VITALS_DIM)),
                  bd.branch(lambda: branch(1, "labs", l, LABS_DIM)))
    c = bd.linear(cat, "fc1.weight", "fc1.bias", 128, 1024, ACT_RELU)
    c = bd.linear(c, "fc2.weight", "fc2.bias", 64, 128, ACT_RELU)
    c = bd.linear(c, "fc3.weight", "fc3.bias", 32, 64, ACT_RELU)
    logit = bd.linear(c, "output.weight", "output.bias", 1, 32)
    bd.bce(logit, y)


DEFAULT_TILE = {"TransformerModel": 16, "RNNModel": 16, "CNNModel": 4}
SUPPORTED = tuple(DEFAULT_TILE)


def build_program(model_name: str, layout: FlatLayout, S: Optional[int] = None, train: bool = True,
                  dropout: bool = True, module: Optional[torch.nn.Module] = None, chain: bool = False) -> Program:
    """Program for one of the ICU models; ``module`` supplies non-default widths (scaled Transformer)."""
    if model_name not in SUPPORTED:
        raise ValueError(f"no native program for model '{model_name}'")
    S = S or DEFAULT_TILE[model_name]
    bd = Builder(layout, S, train, reuse_grads=not chain)
    use_dropout = dropout and train
    if model_name == "TransformerModel":
        d = layout.by_name["vitals_dense.weight"].shape[0]
        ff = layout.by_name["vitals_transformer.ffn.0.weight"].shape[0]
        heads = module.vitals_transformer.attention.num_heads if module is not None else 4
        _transformer(bd, d, ff, heads, use_dropout)
    elif model_name == "RNNModel":
        H = layout.by_name["vitals_gru1.weight_hh_l0"].shape[1]
        _rnn(bd, H, use_dropout)
    else:
        _cnn(bd, use_dropout)
    prog = bd.finish(model_name)
    prog.chain = bool(chain and train)
    return prog
