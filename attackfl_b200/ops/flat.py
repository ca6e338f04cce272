"""Python bindings of the flat-vector kernels (K1/K2 FedAvg+moments, K4 craft, K6 select, T8 optimisers,
K5a/K5b hyper-network).  Every function dispatches on the tensor's device: CUDA tensors go to the sm_100a
library, CPU tensors to the CPU twin built from the same sources."""
from __future__ import annotations

import ctypes
from typing import List, Optional, Sequence

import torch


from .native import check, lib_for, ptr, stream_ptr

c_i, c_f, c_i64, c_u64, c_u32 = ctypes.c_int, ctypes.c_float, ctypes.c_int64, ctypes.c_uint64, ctypes.c_uint32
MAX_ATT = 4
BETAS = (0.9, 0.999)
EPS = 1e-8


def adam_flat(w, g, m, v, lr: float, step: int, betas=BETAS, eps: float = EPS, gscale: Optional[torch.Tensor] = None):
    """In-place Adam on flat fp32 tensors (any shape, numel % 4 == 0); ``step`` is 1-based."""
    lib = lib_for(w)
    rc = lib.afl_adam_flat(ptr(w), ptr(g), ptr(m), ptr(v), c_i64(w.numel()), c_f(lr), c_f(betas[0]), c_f(betas[1]),
                           c_f(eps), c_i(step), ptr(gscale), stream_ptr(w))
    check(lib, rc, "afl_adam_flat")


def sgdm_flat(w, g, buf, lr: float, momentum: float, first_step: bool, gscale: Optional[torch.Tensor] = None):
    lib = lib_for(w)
    rc = lib.afl_sgdm_flat(ptr(w), ptr(g), ptr(buf), c_i64(w.numel()), c_f(lr), c_f(momentum),
                           c_i(1 if first_step else 0), ptr(gscale), stream_ptr(w))
    check(lib, rc, "afl_sgdm_flat")


def fedavg(rows: torch.Tensor, weights: torch.Tensor, out: Optional[torch.Tensor] = None,
           bcast: Optional[torch.Tensor] = None, *, member: Optional[torch.Tensor] = None,
           z: Optional[torch.Tensor] = None, ref: Optional[torch.Tensor] = None, want_mu: bool = False,
           want_sd: bool = False, want_lie: bool = False, strict_k1: bool = False):
    """K1 (+K2).  ``rows [C, d_pad]``, ``weights [C]`` (already normalised).  ``bcast [Cb, d_pad]`` receives the
    result in every row (the START fan-out).  ``member [A, C]`` (0/1 floats) selects each attacker's genuine
    subset; returns ``(out, mu, sd, lie)`` with ``mu/sd/lie`` of shape ``[A, d_pad]`` or ``None``."""
    C, d_pad = rows.shape
    dev = rows.device
    lib = lib_for(rows)
    if out is None:
        out = torch.empty(d_pad, dtype=torch.float32, device=dev)
    A = 0 if member is None else member.shape[0]
    mu = torch.empty(A, d_pad, dtype=torch.float32, device=dev) if (A and (want_mu or want_lie)) else None
    sd = torch.empty(A, d_pad, dtype=torch.float32, device=dev) if (A and (want_sd or want_lie)) else None
    lie = torch.empty(A, d_pad, dtype=torch.float32, device=dev) if (A and want_lie) else None
    weights = weights.to(device=dev, dtype=torch.float32).contiguous()
    if A == 0:
        rc = lib.afl_fedavg(ptr(rows), ptr(weights), c_i(C), c_i64(d_pad), ptr(out), ptr(bcast),
                            c_i(0 if bcast is None else bcast.shape[0]), c_i(0), ptr(None), ptr(None), ptr(None),
                            ptr(None), ptr(None), ptr(None), ptr(None), c_i(0), stream_ptr(rows))
        check(lib, rc, "afl_fedavg")
        return out, None, None, None
    member = member.to(device=dev, dtype=torch.float32).contiguous()
    kcount = member.sum(1).contiguous()
    z = (torch.zeros(A, device=dev) if z is None else z.to(device=dev, dtype=torch.float32)).contiguous()
    for a0 in range(0, A, MAX_ATT):
        a1 = min(A, a0 + MAX_ATT)
        first = a0 == 0
        rc = lib.afl_fedavg(ptr(rows), ptr(weights), c_i(C), c_i64(d_pad), ptr(out) if first else ptr(None),
                            ptr(bcast) if first else ptr(None), c_i(0 if (bcast is None or not first) else bcast.shape[0]),
                            c_i(a1 - a0), ptr(ref), ptr(member[a0:a1]), ptr(kcount[a0:a1]), ptr(z[a0:a1]),
                            ptr(mu[a0:a1]) if mu is not None else ptr(None),
                            ptr(sd[a0:a1]) if sd is not None else ptr(None),
                            ptr(lie[a0:a1]) if lie is not None else ptr(None), c_i(1 if strict_k1 else 0),
                            stream_ptr(rows))
        check(lib, rc, "afl_fedavg")
    return out, mu, sd, lie


def gamma_search(G: torch.Tensor, k: int, rule: str, exclude_first: bool = True, paper: bool = False) -> torch.Tensor:
    """K4 bisection on a ``(k+2)x(k+2)`` fp32 Gram matrix; returns device tensor ``[gamma, gamma_succ, thr, probes]``."""
    lib = lib_for(G)
    G = G.to(torch.float32).contiguous()
    out = torch.zeros(4, dtype=torch.float32, device=G.device)
    rc = lib.afl_gamma_search_dev(ptr(G), c_i(k), c_i(0 if rule == "max" else 1), c_i(1 if exclude_first else 0),
                                  c_i(1 if paper else 0), ptr(out), stream_ptr(G))
    check(lib, rc, "afl_gamma_search_dev")
    return out


def craft_axpy(mu: torch.Tensor, p: Optional[torch.Tensor], gamma: torch.Tensor, sign_mode: bool = False,
               out: Optional[torch.Tensor] = None) -> torch.Tensor:
    lib = lib_for(mu)
    out = torch.empty_like(mu) if out is None else out
    rc = lib.afl_craft_axpy(ptr(out), ptr(mu), ptr(p), ptr(gamma), c_i(1 if sign_mode else 0), c_i64(mu.numel()),
                            stream_ptr(mu))
    check(lib, rc, "afl_craft_axpy")
    return out


def craft_random(w: torch.Tensor, valid: Optional[torch.Tensor], perturbation: float, seed: int, stream_id: int,
                 out: Optional[torch.Tensor] = None) -> torch.Tensor:
    lib = lib_for(w)
    out = torch.empty_like(w) if out is None else out
    rc = lib.afl_craft_random(ptr(out), ptr(w), ptr(valid), c_f(perturbation), c_u64(seed & (2 ** 64 - 1)),
                              c_u32(stream_id), c_i64(w.numel()), stream_ptr(w))
    check(lib, rc, "afl_craft_random")
    return out


def coord_select(rows: torch.Tensor, mode: str, trim: int = 0, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """K6: ``mode='median'`` (lower median) or ``'trimmed'`` (drop ``trim`` per side, mean of the rest)."""
    C, d_pad = rows.shape
    lib = lib_for(rows)
    out = torch.empty(d_pad, dtype=torch.float32, device=rows.device) if out is None else out
    rc = lib.afl_coord_select(ptr(rows), c_i(C), c_i64(d_pad), ptr(out), c_i(0 if mode == "median" else 1),
                              c_i(trim), stream_ptr(rows))
    check(lib, rc, "afl_coord_select")
    return out


ROC_AUC_MAX = 8192


def roc_auc(scores: torch.Tensor, labels: torch.Tensor) -> torch.Tensor:
    """``[auc, #NaN scores, #positives, #negatives]`` (device tensor, no host sync) for up to ``ROC_AUC_MAX`` scores:
    the tie-aware trapezoid of ``validation.roc_auc`` in one launch."""
    s = scores.reshape(-1).to(torch.float32).contiguous()
    y = labels.reshape(-1).to(device=s.device, dtype=torch.float32).contiguous()
    assert s.numel() == y.numel() and 1 <= s.numel() <= ROC_AUC_MAX
    lib = lib_for(s)
    out = torch.empty(4, dtype=torch.float32, device=s.device)
    check(lib, lib.afl_roc_auc(ptr(s), ptr(y), c_i64(s.numel()), ptr(out), stream_ptr(s)), "afl_roc_auc")
    return out


def round_indices(n_rows: int, sizes: torch.Tensor, client_ids: torch.Tensor, epochs: int, max_size: int, seed: int) -> torch.Tensor:
    """Per-round sample indices ``[C, epochs, max_size]`` int32 (-1 past a client's size): a random subset of ``sizes[c]``
    of the ``n_rows`` records per client and a fresh order per epoch (``random.sample`` + ``DataLoader(shuffle=True)``,
    ``src/RpcClient.py:166,169``), as keyed permutations evaluated per element — one launch, no sort, the same indices on
    the CPU twin and on the GPU.  ``sizes`` / ``client_ids``: int32 tensors on the target device."""
    assert sizes.dtype == torch.int32 and client_ids.dtype == torch.int32 and sizes.numel() == client_ids.numel()
    C = sizes.numel()
    lib = lib_for(sizes)
    idx = torch.empty(C, epochs, max_size, dtype=torch.int32, device=sizes.device)
    check(lib, lib.afl_round_indices(ptr(idx), ptr(sizes), ptr(client_ids), c_i(C), c_i(epochs), c_i(max_size), c_i(n_rows),
                                     c_u64(seed & (2 ** 64 - 1)), stream_ptr(sizes)), "afl_round_indices")
    return idx


def scion_quantize(rows: torch.Tensor, valid: Optional[torch.Tensor], seed: int, row0: int = 0,
                   bits: Optional[torch.Tensor] = None):
    """ScionFL 1-bit stochastic quantisation of every row (``src/Utils.py:372-376``), packed 32 coordinates per word.

    Returns ``(bits [C, d_pad/32] int32, counts [C] int32, minmax [C, 2])``; ``valid`` is the layout's 0/1 float mask.
    ``row0``: Philox stream id of row 0 (N GPUs: the position of this rank's first row in the round's reporting list, so
    that the bits equal those of the one-GPU pass over all rows); ``bits``: optional output buffer (symmetric memory)."""
    C, d_pad = rows.shape
    lib = lib_for(rows)
    dev = rows.device
    minmax = torch.empty(C, 2, dtype=torch.float32, device=dev)
    bits = torch.empty(C, d_pad // 32, dtype=torch.int32, device=dev) if bits is None else bits
    assert bits.shape[0] >= C and bits.shape[1] == d_pad // 32 and bits.is_contiguous()
    counts = torch.empty(C, dtype=torch.int32, device=dev)
    check(lib, lib.afl_scion_minmax(ptr(rows), ptr(valid), c_i(C), c_i64(d_pad), ptr(minmax), stream_ptr(rows)),
          "afl_scion_minmax")
    check(lib, lib.afl_scion_quantize_rows(ptr(rows), ptr(valid), ptr(minmax), c_i(C), c_i64(d_pad), c_u64(seed),
                                           c_i(int(row0)), ptr(bits), ptr(counts), stream_ptr(rows)), "afl_scion_quantize")
    return bits[:C], counts, minmax


def scion_scores_ptrs(bits_ptrs: Sequence[int], words: int, w_lo: int, w_hi: int, valid: Optional[torch.Tensor],
                      lohi: torch.Tensor) -> torch.Tensor:
    """``scion_scores`` restricted to the words ``[w_lo, w_hi)`` with every client's packed row given by address
    (peer-mapped addresses allowed): the partial sums of a slice owner."""
    C = len(bits_ptrs)
    dev = lohi.device
    lib = lib_for(lohi)
    out = torch.empty(C + 2, dtype=torch.float32, device=dev)
    lohi = lohi.to(torch.float32).contiguous()
    bp = torch.as_tensor([int(p) for p in bits_ptrs], dtype=torch.int64, device=dev)
    check(lib, lib.afl_scion_scores_ptrs(ptr(bp), ptr(valid), ptr(lohi), c_i(C), c_i64(words * 32), c_i64(w_lo), c_i64(w_hi),
                                         ptr(out), stream_ptr(lohi)), "afl_scion_scores_ptrs")
    return out


def scion_distances_from_sums(sums: torch.Tensor, counts: torch.Tensor, lo: torch.Tensor, hi: torch.Tensor, d: float):
    """Cosine distances from ``sums = [S_0..S_{C-1}, sum agg, sum agg^2]`` and the (clipped) levels (server.py:455-466)."""
    C = counts.numel()
    out = sums.to(torch.float64)
    s, a_sum, a_sq = out[:C], out[C], out[C + 1]
    c1 = counts.to(torch.float64)
    dots = lo * a_sum + (hi - lo) * s
    n_deq = torch.sqrt((d - c1) * lo * lo + c1 * hi * hi)
    cos = dots / (n_deq.clamp_min(1e-8) * torch.sqrt(a_sq).clamp_min(1e-8))
    return (1.0 - cos).to(torch.float32)


def scion_clipped_levels(counts: torch.Tensor, minmax: torch.Tensor, d: float, mu_threshold: float = 3.0):
    """L2 norms from the bit counts and the levels after the ``mu_threshold`` x mean clip (server.py:443-454)."""
    c1 = counts.to(torch.float64)
    lo, hi = minmax[:, 0].to(torch.float64), minmax[:, 1].to(torch.float64)
    l2 = torch.sqrt((d - c1) * lo * lo + c1 * hi * hi)
    cap = mu_threshold * l2.mean()
    fct = torch.where(l2 > cap, cap / l2, torch.ones_like(l2))
    return lo * fct, hi * fct


def scion_scores(bits: torch.Tensor, valid: Optional[torch.Tensor], lohi: torch.Tensor) -> torch.Tensor:
    """``out[:C] = sum_j sigma_cj agg_j``, ``out[C] = sum_j agg_j``, ``out[C+1] = sum_j agg_j^2`` for the mean ``agg`` of
    the dequantised rows ``lo_c + sigma_cj (hi_c - lo_c)`` — computed from the packed bits only."""
    C, words = bits.shape
    lib = lib_for(bits)
    out = torch.empty(C + 2, dtype=torch.float32, device=bits.device)
    lohi = lohi.to(torch.float32).contiguous()
    check(lib, lib.afl_scion_scores(ptr(bits), ptr(valid), ptr(lohi), c_i(C), c_i64(words * 32), ptr(out),
                                    stream_ptr(bits)), "afl_scion_scores")
    return out


def scion_cosine_distances(rows: torch.Tensor, valid: Optional[torch.Tensor], seed: int, mu_threshold: float = 3.0):
    """The whole ScionFL scoring pass (``server.py:443-466``): quantise, L2 from bit counts, clip at
    ``mu_threshold`` x mean, cosine distance of every dequantised row to their mean.  Returns ``(cosd [C], aux)``."""
    C, d_pad = rows.shape
    bits, counts, minmax = scion_quantize(rows, valid, seed)
    d = float(valid.sum().item()) if valid is not None else float(d_pad)
    lo, hi = scion_clipped_levels(counts, minmax, d, mu_threshold)
    out = scion_scores(bits, valid, torch.stack([lo, hi], dim=1))
    cosd = scion_distances_from_sums(out, counts, lo, hi, d)
    return cosd, {"bits": bits, "counts": counts, "minmax": minmax, "lohi": torch.stack([lo, hi], 1)}


def coord_select_ptrs(row_ptrs: Sequence[int], lo: int, hi: int, dst_ptrs: Sequence[int], mode: str, trim: int,
                      device) -> None:
    """K6 on a coordinate slice with the rows given by device addresses (peer-mapped addresses allowed); the result is
    stored into every buffer of ``dst_ptrs``."""
    lib = lib_for(torch.empty(0, device=device))
    rp = torch.as_tensor([int(p) for p in row_ptrs], dtype=torch.int64, device=device)
    dp = torch.as_tensor([int(p) for p in dst_ptrs], dtype=torch.int64, device=device)
    rc = lib.afl_coord_select_ptrs(ptr(rp), c_i(len(row_ptrs)), c_i64(lo), c_i64(hi), ptr(dp), c_i(len(dst_ptrs)),
                                   c_i(0 if mode == "median" else 1), c_i(trim), stream_ptr(rp))
    check(lib, rc, "afl_coord_select_ptrs")


def sliced_gram(row_ptrs: Sequence[int], center: torch.Tensor, lo: int, hi: int, precision: str = "tf32",
                reduce_gram=None) -> torch.Tensor:
    """Gram of ``rows - center`` from the coordinate slice ``[lo, hi)`` of rows given by device addresses (read in place,
    also over NVLink); ``reduce_gram`` sums the partial Grams of the ranks.  Serves Krum / ShieldFL / FLTrust."""
    from . import gram as gram_mod
    dev = center.device
    lib = lib_for(center)
    n, ld = len(row_ptrs), hi - lo
    rp = torch.as_tensor([int(p) for p in row_ptrs], dtype=torch.int64, device=dev)
    X = torch.empty(n + 1, ld, dtype=torch.float32, device=dev)
    mu = torch.empty(1, ld, dtype=torch.float32, device=dev)
    p = torch.empty(1, ld, dtype=torch.float32, device=dev)
    one = torch.ones(1, n, dtype=torch.float32, device=dev)
    rc = lib.afl_multi_moments(ptr(rp), c_i(n), ptr(center), c_i64(lo), c_i64(hi), c_i64(ld), ptr(one),
                               ptr(torch.tensor([float(n)], device=dev)), ptr(torch.zeros(1, dtype=torch.int32, device=dev)),
                               c_i(1), ptr(X), ptr(mu), ptr(p), stream_ptr(center))
    check(lib, rc, "afl_multi_moments")
    G = gram_mod.gram(X[:n], precision=precision if n <= 256 else "fp32")
    if reduce_gram is not None:
        G = reduce_gram(G)
    return G


def multi_gram_attack(row_ptrs: Sequence[int], center: torch.Tensor, subsets: Sequence[Sequence[int]], modes: Sequence[str],
                      d_pad: int, lo: int = 0, hi: Optional[int] = None, paper: bool = False, precision: str = "tf32",
                      dst_ptrs: Optional[Sequence[Sequence[int]]] = None, reduce_gram=None):
    """Min-Max / Min-Sum / Opt-Fang for ALL attackers from one Gram (``csrc/gram_multi.cu``).

    ``row_ptrs``: device addresses of the n genuine rows (peer-mapped addresses allowed); ``subsets[a]``: positions in
    that list, in the attacker's sample order; ``modes[a]``; slice ``[lo, hi)`` of the coordinates handled by this call
    (``reduce_gram`` sums the R x R partial Grams across ranks when the slices are distributed).  Crafted slices are
    written through ``dst_ptrs[a][w]`` (every rank's copy of attacker ``a``'s row) or returned as ``[A, hi-lo]``.
    Returns ``(crafted or None, gammas [A,4] device tensor)``.
    """
    from . import gram as gram_mod
    dev = center.device
    lib = lib_for(center)
    hi = d_pad if hi is None else hi
    n, A = len(row_ptrs), len(subsets)
    ld = hi - lo
    kmax = max(len(s) for s in subsets)
    rp = torch.as_tensor([int(p) for p in row_ptrs], dtype=torch.int64, device=dev)
    member = torch.zeros(A, n, dtype=torch.float32)
    sub = torch.zeros(A, kmax, dtype=torch.int32)
    for a, s in enumerate(subsets):
        member[a, list(s)] = 1.0
        sub[a, :len(s)] = torch.as_tensor(list(s), dtype=torch.int32)
    member, sub = member.to(dev), sub.to(dev)
    kc = torch.tensor([float(len(s)) for s in subsets], device=dev)
    kk = torch.tensor([len(s) for s in subsets], dtype=torch.int32, device=dev)
    sgn = torch.tensor([1 if m == "Opt-Fang" else 0 for m in modes], dtype=torch.int32, device=dev)
    rule = torch.tensor([1 if m == "Min-Sum" else 0 for m in modes], dtype=torch.int32, device=dev)
    X = torch.empty(n + A, ld, dtype=torch.float32, device=dev)
    mu = torch.empty(A, ld, dtype=torch.float32, device=dev)
    p = torch.empty(A, ld, dtype=torch.float32, device=dev)
    rc = lib.afl_multi_moments(ptr(rp), c_i(n), ptr(center), c_i64(lo), c_i64(hi), c_i64(ld), ptr(member), ptr(kc), ptr(sgn),
                               c_i(A), ptr(X), ptr(mu), ptr(p), stream_ptr(center))
    check(lib, rc, "afl_multi_moments")
    G = gram_mod.gram(X, precision=precision)
    if reduce_gram is not None:
        G = reduce_gram(G)
    G = G.to(torch.float32).contiguous()
    gam = torch.zeros(A, 4, dtype=torch.float32, device=dev)
    rc = lib.afl_multi_gamma(ptr(G), c_i(n + A), c_i(n), ptr(sub), ptr(kk), c_i(kmax), ptr(rule), c_i(A),
                             c_i(1 if paper else 0), ptr(gam), stream_ptr(center))
    check(lib, rc, "afl_multi_gamma")
    out = None
    if dst_ptrs is None:
        out = torch.empty(A, ld, dtype=torch.float32, device=dev)
        dst = torch.as_tensor([out[a].data_ptr() - lo * 4 for a in range(A)], dtype=torch.int64, device=dev)
        world = 1
    else:
        world = len(dst_ptrs[0])
        dst = torch.as_tensor([int(x) for row in dst_ptrs for x in row], dtype=torch.int64, device=dev)
    rc = lib.afl_multi_craft(ptr(dst), c_i(world), ptr(mu), ptr(p), c_i64(ld), ptr(gam), c_i(A), c_i64(lo), c_i64(hi),
                             stream_ptr(center))
    check(lib, rc, "afl_multi_craft")
    return out, gam


# ---------------------------------------------------------------------------------------- hyper-network
class NativeHyper:
    """Device-side state of a ``FusedHyperNet`` in the layout the K5 kernels expect.

    ``small = params | m | v`` where ``params = emb | w0 b0 | w1 b1 | w2 b2``; the head lives in
    ``head_w [D_pad, H]``, ``head_b [D_pad]`` with Adam moments ``mw, vw, mb, vb``.
    """

    def __init__(self, fused):
        from ..models.hyper import FusedHyperNet
        assert isinstance(fused, FusedHyperNet) and fused.n_hidden == 2
        self.fused = fused
        dev = fused.device
        self.n_nodes, self.E, self.H = fused.n_nodes, fused.E, fused.H
        self.tot = self.n_nodes * self.E + self.H * self.E + self.H + 2 * (self.H * self.H + self.H)
        self.small = torch.zeros(3 * self.tot, dtype=torch.float32, device=dev)
        self._bind_head()
        self.delta = torch.zeros(fused.layout.padded, dtype=torch.float32, device=dev)
        self.acts = torch.zeros(self.E + 5 * self.H, dtype=torch.float32, device=dev)
        self.red = torch.zeros(self.H + 3, dtype=torch.float32, device=dev)
        self.f1 = torch.zeros(self.H, dtype=torch.float32, device=dev)
        self._client_idx = torch.arange(self.n_nodes, dtype=torch.int32, device=dev)
        self.pack()

    def _bind_head(self) -> None:
        fused = self.fused
        self.head_w, self.head_b = fused.head_w, fused.head_b
        self.mw, self.vw = fused.m["head_w"], fused.v["head_w"]
        self.mb, self.vb = fused.m["head_b"], fused.v["head_b"]

    def _names(self) -> List[str]:
        return ["emb", "mlp_w0", "mlp_b0", "mlp_w1", "mlp_b1", "mlp_w2", "mlp_b2"]

    def pack(self) -> None:
        """fused tensors -> ``small``."""
        for base, src in ((0, self.fused.trainable()), (self.tot, self.fused.m), (2 * self.tot, self.fused.v)):
            o = base
            for n in self._names():
                t = src[n].reshape(-1)
                self.small[o:o + t.numel()].copy_(t)
                o += t.numel()

    def unpack(self) -> None:
        """``small`` -> fused tensors (so export / torch paths see the native updates)."""
        for base, dst in ((0, self.fused.trainable()), (self.tot, self.fused.m), (2 * self.tot, self.fused.v)):
            o = base
            for n in self._names():
                t = dst[n]
                t.copy_(self.small[o:o + t.numel()].view_as(t))
                o += t.numel()

    def emit(self, clients: Sequence[int], out: Optional[torch.Tensor] = None, chunk: int = 64) -> torch.Tensor:
        """K5a: rows ``theta[j] = h(clients[j])``; ``out [n, D_pad]`` may be a slice of the client arena."""
        self.fused._need_full("NativeHyper.emit")
        dev = self.small.device
        lib = lib_for(self.small)
        d_pad = self.fused.layout.padded
        n = len(clients)
        out = torch.empty(n, d_pad, dtype=torch.float32, device=dev) if out is None else out
        cl = torch.as_tensor(list(clients), dtype=torch.int32, device=dev)
        f = torch.empty(max(n, 1) * self.H, dtype=torch.float32, device=dev)
        if dev.type == "cuda":
            # same arithmetic as pass A of the round kernel (csrc/hyper_round.cu): a coordinate the client did not change
            # then has delta == 0 exactly, as in the reference (theta - theta_client of identical tensors), instead of
            # rounding noise that Adam would normalise to +-lr
            tp = torch.as_tensor([out[j].data_ptr() for j in range(n)], dtype=torch.int64, device=dev)
            for c0 in range(0, n, chunk):
                c1 = min(n, c0 + chunk)
                rc = lib.afl_hyper_emit_rows(ptr(self.head_w), ptr(self.head_b), c_i64(d_pad), c_i64(0), ptr(self.small),
                                             c_i(self.n_nodes), c_i(self.E), c_i(self.H), ptr(cl[c0:c1]), c_i(c1 - c0),
                                             ptr(f), ptr(tp[c0:c1]), stream_ptr(self.small))
                check(lib, rc, "afl_hyper_emit_rows")
            return out
        for c0 in range(0, n, chunk):
            c1 = min(n, c0 + chunk)
            rc = lib.afl_hyper_emit(ptr(self.head_w), ptr(self.head_b), ptr(self.small), c_i(self.n_nodes), c_i(self.E),
                                    c_i(self.H), ptr(cl[c0:c1]), c_i(c1 - c0), ptr(f), ptr(self.acts),
                                    ptr(out[c0:c1]), c_i64(out.stride(0)), c_i64(d_pad), stream_ptr(self.small))
            check(lib, rc, "afl_hyper_emit")
        return out

    def step(self, client: int, theta_client: torch.Tensor, lr: float, clip: float) -> None:
        """K5b: one sequential pFedHN step (no host sync; the clip coefficient stays on the device)."""
        self.fused._need_full("NativeHyper.step")
        lib = lib_for(self.small)
        self.fused.step += 1
        rc = lib.afl_hyper_step(ptr(self.head_w), ptr(self.head_b), ptr(self.mw), ptr(self.vw), ptr(self.mb),
                                ptr(self.vb), ptr(self.small), c_i(self.n_nodes), c_i(self.E), c_i(self.H),
                                c_i(int(client)), ptr(self._client_idx[int(client):int(client) + 1]), ptr(theta_client),
                                ptr(self.delta), ptr(self.f1), ptr(self.acts), ptr(self.red),
                                c_i64(self.fused.layout.padded), c_f(lr), c_f(BETAS[0]), c_f(BETAS[1]), c_f(EPS),
                                c_i(self.fused.step), c_f(clip), stream_ptr(self.small))
        check(lib, rc, "afl_hyper_step")

    def embedding(self, i: int) -> torch.Tensor:
        return self.small[i * self.E:(i + 1) * self.E].view(1, self.E).clone()

    # ------------------------------------------------------------------ persistent / row-sharded round (csrc/hyper_round.cu)
    def setup_round(self, lo: int = 0, hi: Optional[int] = None, world: int = 1, rank: int = 0, n_max: int = 0,
                    xchg_ptrs=None, sig_ptrs=None, shard_storage: bool = False) -> None:
        """Row shard ``[lo, hi)`` of the head owned by this rank (``world`` ranks) and the symmetric exchange / flag
        buffers (host arrays of peer-mapped pointers) for up to ``n_max`` clients per round.  ``shard_storage``: also
        release the rows this rank does not own (``FusedHyperNet.shard_rows``): W_h, b_h and both Adam moments then take
        ``1 / world`` of their size on every GPU."""
        dev = self.small.device
        hi = self.fused.layout.padded if hi is None else hi
        if shard_storage and not self.fused.sharded and (lo, hi) != (0, self.fused.layout.padded):
            self.fused.shard_rows(int(lo), int(hi))
            self._bind_head()
            self.delta = None                      # full-length scratch of the one-GPU step: not used any more
        self.lo, self.hi, self.world, self.rank = int(lo), int(hi), int(world), int(rank)
        self.n_max = int(n_max) if n_max else self.n_nodes
        self.xchg_ptrs, self.sig_ptrs = xchg_ptrs, sig_ptrs
        self.small_alt = torch.zeros(self.tot, dtype=torch.float32, device=dev)
        self.delta_s = torch.zeros(self.hi - self.lo, dtype=torch.float32, device=dev)
        self.red2 = torch.zeros(2 * (self.H + 1), dtype=torch.float32, device=dev)
        self.bar = torch.zeros(2, dtype=torch.int32, device=dev)
        self.norms = torch.zeros(self.n_max, dtype=torch.float32, device=dev)
        self.f_all = torch.zeros(self.n_max * self.H, dtype=torch.float32, device=dev)
        self.round_epoch = 0

    def _shards(self):
        if self.fused.sharded:                     # the tensors ARE the shard
            assert (self.fused.row_lo, self.fused.row_hi) == (self.lo, self.hi)
            return (self.head_w, self.head_b, self.mw, self.vw, self.mb, self.vb)
        lo, hi = self.lo, self.hi
        return (self.head_w[lo:hi], self.head_b[lo:hi], self.mw[lo:hi], self.vw[lo:hi], self.mb[lo:hi], self.vb[lo:hi])

    def round(self, order: Sequence[int], theta_ptrs: Sequence[int], lr: float, clip: float) -> torch.Tensor:
        """All sequential pFedHN steps of one round (``server.py:644-670``) in ONE launch on this rank's row shard.
        ``theta_ptrs[j]``: device address of the full row of client ``order[j]`` (peer-mapped on other ranks).
        Returns the pre-clip gradient norms (device tensor, no host sync)."""
        if getattr(self, "lo", None) is None:
            self.setup_round()
        lib = lib_for(self.small)
        dev = self.small.device
        n = len(order)
        assert n <= self.n_max
        od = torch.as_tensor(list(order), dtype=torch.int32, device=dev)
        tp = torch.as_tensor([int(p) for p in theta_ptrs], dtype=torch.int64, device=dev)
        hw, hb, mw, vw, mb, vb = self._shards()
        self.round_epoch += 1
        rc = lib.afl_hyper_round(ptr(hw), ptr(hb), ptr(mw), ptr(vw), ptr(mb), ptr(vb), c_i64(self.hi - self.lo),
                                 c_i64(self.lo), ptr(self.small), ptr(self.small_alt), c_i(self.n_nodes), c_i(self.E),
                                 c_i(self.H), ptr(od), ptr(tp), c_i(n), ptr(self.delta_s), ptr(self.red2),
                                 self.xchg_ptrs, self.sig_ptrs, c_i(self.world), c_i(self.rank), c_i(self.n_max),
                                 ctypes.c_uint32(self.round_epoch), ptr(self.bar), c_i(self.fused.step), c_f(lr),
                                 c_f(BETAS[0]), c_f(BETAS[1]), c_f(EPS), c_f(clip), ptr(self.norms), stream_ptr(self.small))
        check(lib, rc, "afl_hyper_round")
        self.fused.step += n
        return self.norms[:n]

    def emit_rows(self, clients: Sequence[int], theta_ptrs: Sequence[int], chunk: int = 64) -> None:
        """K5a scatter: this rank's row shard of ``theta_j = h(clients[j])`` stored straight into the clients' rows."""
        if getattr(self, "lo", None) is None:
            self.setup_round()
        lib = lib_for(self.small)
        dev = self.small.device
        hw, hb = self._shards()[:2]
        cl = torch.as_tensor(list(clients), dtype=torch.int32, device=dev)
        tp = torch.as_tensor([int(p) for p in theta_ptrs], dtype=torch.int64, device=dev)
        for c0 in range(0, len(clients), chunk):
            c1 = min(len(clients), c0 + chunk)
            if self.f_all.numel() < (c1 - c0) * self.H:
                self.f_all = torch.zeros((c1 - c0) * self.H, dtype=torch.float32, device=dev)
            rc = lib.afl_hyper_emit_rows(ptr(hw), ptr(hb), c_i64(self.hi - self.lo), c_i64(self.lo), ptr(self.small),
                                         c_i(self.n_nodes), c_i(self.E), c_i(self.H), ptr(cl[c0:c1]), c_i(c1 - c0),
                                         ptr(self.f_all), ptr(tp[c0:c1]), stream_ptr(self.small))
            check(lib, rc, "afl_hyper_emit_rows")

    def check_round_error(self) -> int:
        return int(self.bar[1].item())
