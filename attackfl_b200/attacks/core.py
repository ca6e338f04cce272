"""Model-poisoning attacks on flat rows: Random, LIE, Min-Max, Min-Sum, Opt-Fang.

Reference parity: ``/root/reference/src/Utils.py:52-57`` (Random), ``:83-98`` (mean/std/sign),
``:101-132`` (Opt-Fang), ``:135-166`` (Min-Max), ``:169-204`` (Min-Sum), ``:207-214`` (LIE),
``:30-49`` (``compute_distance``) and the dispatch at ``src/RpcClient.py:119-145``.

The reference rebuilds a full model and walks every tensor for each of the 6 gamma probes and for
all k(k-1)/2 pairs (9-18 s at k=48 on CPU).  Here the whole search runs on the
``(k+2) x (k+2)`` Gram matrix of ``[g_1-c .. g_k-c, mu-c, p]``: one pass over D, then
O(k) scalars per probe.  ``gram_fn`` is pluggable: the engine passes the tcgen05 split-K kernel
(``ops.gram``), the default is an fp32 torch matmul.

Quirks kept under ``search='reference'``: gamma0=50, tau=1; the *last built* model is
returned, not m(gamma_succ); genuine model 0 is excluded from the acceptance test (aliasing at
``src/Utils.py:120,154,192``); ``len(S) <= 1`` returns the attacker's own weights.
Metric modes: ``flat`` (one L2 over the flat vector, paper-faithful, Gram),
``layer-fro`` (sum over tensors of Frobenius norms, segmented Gram), ``reference`` (exact
per-tensor ``torch.linalg.norm(ord=2)``: spectral norm for 2-D tensors; oracle only, CPU-speed).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import Callable, List, Optional, Tuple

import torch

from ..layout import FlatLayout

GAMMA0 = 50.0
TAU = 1.0

GramFn = Callable[[torch.Tensor], torch.Tensor]


def _default_gram(a: torch.Tensor) -> torch.Tensor:
    a = a.to(torch.float32)
    return a @ a.t()


# ---------------------------------------------------------------------------- statistics
def mean_std(rows: torch.Tensor, strict_reference: bool = False) -> Tuple[torch.Tensor, torch.Tensor]:
    """Per-coordinate mean and *unbiased* std (``torch.std`` default, Utils.py:89-90).

    k=1 gives NaN std in the reference; unless ``strict_reference`` we return 0.
    Uses sums shifted by the first row (cancellation-safe, same as the fused K2 kernel).
    """
    k = rows.shape[0]
    ref = rows[0:1]
    d = rows - ref
    s1 = d.sum(0)
    mean = ref[0] + s1 / k
    if k < 2:
        std = torch.full_like(mean, float("nan")) if strict_reference else torch.zeros_like(mean)
        return mean, std
    s2 = (d * d).sum(0)
    var = (s2 - s1 * s1 / k) / (k - 1)
    return mean, var.clamp_min(0).sqrt()


def random_attack(own: torch.Tensor, perturbation: float, valid: Optional[torch.Tensor] = None,
                  generator: Optional[torch.Generator] = None) -> torch.Tensor:
    """``own + N(0,1) * perturbation`` (Utils.py:52-57); padding coordinates stay zero."""
    noise = torch.randn(own.shape, device=own.device, dtype=own.dtype, generator=generator)
    out = own + noise * float(perturbation)
    if valid is not None:
        out = out * valid.to(out.dtype)
    return out


def lie_attack(rows: torch.Tensor, z: float, strict_reference: bool = False) -> torch.Tensor:
    """``mu + z * sigma`` (Utils.py:207-214)."""
    mu, sd = mean_std(rows, strict_reference)
    return mu + float(z) * sd


# ---------------------------------------------------------------------------- distances
def _per_tensor_gram(a: torch.Tensor, layout: FlatLayout) -> torch.Tensor:
    """``[T, R, R]`` Gram per tensor segment (layer-fro metric)."""
    out = []
    for e in layout.entries:
        seg = a[:, e.offset:e.offset + e.numel].to(torch.float32)
        out.append(seg @ seg.t())
    return torch.stack(out)


def _dist_from_gram(g: torch.Tensor) -> torch.Tensor:
    """Pairwise distances from a Gram matrix (``[..., R, R]``)."""
    diag = torch.diagonal(g, dim1=-2, dim2=-1)
    d2 = diag.unsqueeze(-1) + diag.unsqueeze(-2) - 2.0 * g
    return d2.clamp_min(0).sqrt()


def reference_distance(a: torch.Tensor, b: torch.Tensor, layout: FlatLayout) -> float:
    """Exact ``compute_distance`` (Utils.py:30-49): sum over tensors of ``linalg.norm(diff, ord=2)``."""
    total = 0.0
    va, vb = layout.views(a), layout.views(b)
    for k in va:
        diff = (va[k] - vb[k]).to("cpu")
        if diff.dim() > 2:
            raise RuntimeError(
                f"attack.metric=reference is undefined for the {diff.dim()}-D tensor '{k}' "
                "(torch.linalg.norm(ord=2) raises in the reference as well); use flat or layer-fro")
        total += torch.linalg.norm(diff, ord=2).item()
    return total


@dataclass
class SearchResult:
    crafted: torch.Tensor        # [D_pad]
    gamma: float                 # gamma of the returned model
    gamma_succ: float            # last accepted gamma (0 if none)
    threshold: float
    probes: List[Tuple[float, bool]]


def gamma_schedule(accept: Callable[[float], bool], gamma0: float = GAMMA0, tau: float = TAU,
                   search: str = "reference") -> Tuple[float, float, List[Tuple[float, bool]]]:
    """The reference's bisection (Utils.py:115-130).  Returns (gamma_returned, gamma_succ, probes)."""
    gamma, step, succ = float(gamma0), float(gamma0), 0.0
    last = gamma
    probes: List[Tuple[float, bool]] = []
    while abs(succ - gamma) > tau:
        last = gamma
        ok = bool(accept(gamma))
        probes.append((gamma, ok))
        if ok:
            succ = gamma
            gamma = gamma + step / 2
        else:
            gamma = gamma - step / 2
        step = step / 2
    if search == "paper":
        return succ, succ, probes
    return last, succ, probes


def _craft_search(own: torch.Tensor, rows: torch.Tensor, direction: str, rule: str,
                  layout: Optional[FlatLayout], metric: str, search: str,
                  center: Optional[torch.Tensor], gram_fn: Optional[GramFn],
                  exclude_first: bool) -> SearchResult:
    k = rows.shape[0]
    if k <= 1:
        return SearchResult(own.clone(), 0.0, 0.0, 0.0, [])
    mu, sd = mean_std(rows)
    p = sd if direction == "std" else torch.sign(mu)
    j0 = 1 if exclude_first else 0

    if metric == "reference":
        assert layout is not None
        dm = torch.zeros(k, k, dtype=torch.float64)
        for i in range(k):
            for j in range(i + 1, k):
                dm[i, j] = dm[j, i] = reference_distance(rows[i], rows[j], layout)
        thr = dm.max().item() if rule == "max" else (dm ** 2).sum(1).max().item()

        def accept(gamma: float) -> bool:
            m = mu - gamma * p
            ds = [reference_distance(m, rows[j], layout) for j in range(j0, k)]
            if rule == "max":
                return max(ds) < thr
            return sum(d * d for d in ds) < thr
    else:
        c = rows.mean(0) if center is None else center
        a = torch.cat([rows - c, (mu - c)[None], p[None]], dim=0)
        if metric == "flat":
            g = (gram_fn or _default_gram)(a).to(torch.float64)[None]       # [1, R, R]
        elif metric == "layer-fro":
            assert layout is not None
            g = _per_tensor_gram(a, layout).to(torch.float64)               # [T, R, R]
        else:
            raise ValueError(f"unknown attack metric '{metric}'")
        gg = g[:, :k, :k]
        pair = _dist_from_gram(gg).sum(0)                                   # [k, k]
        thr = pair.max().item() if rule == "max" else (pair ** 2).sum(1).max().item()
        g_mm = g[:, k, k]
        g_pp = g[:, k + 1, k + 1]
        g_mp = g[:, k, k + 1]
        g_jj = torch.diagonal(gg, dim1=-2, dim2=-1)                         # [T, k]
        g_mj = g[:, k, :k]
        g_pj = g[:, k + 1, :k]

        def accept(gamma: float) -> bool:
            d2 = (g_mm[:, None] + gamma * gamma * g_pp[:, None] + g_jj - 2 * gamma * g_mp[:, None]
                  - 2 * g_mj + 2 * gamma * g_pj)
            d = d2.clamp_min(0).sqrt().sum(0)[j0:]                           # [k - j0]
            if rule == "max":
                return d.max().item() < thr
            return (d * d).sum().item() < thr

    gamma, succ, probes = gamma_schedule(accept, search=search)
    return SearchResult(mu - gamma * p, gamma, succ, float(thr), probes)


def min_max_attack(own, rows, layout=None, metric="flat", search="reference", center=None,
                   gram_fn=None, exclude_first=True) -> SearchResult:
    """``mu - gamma*sigma`` with ``max_j d(m,g_j) < max_{i<j} d(g_i,g_j)`` (Utils.py:135-166)."""
    return _craft_search(own, rows, "std", "max", layout, metric, search, center, gram_fn, exclude_first)


def min_sum_attack(own, rows, layout=None, metric="flat", search="reference", center=None,
                   gram_fn=None, exclude_first=True) -> SearchResult:
    """``mu - gamma*sigma`` with ``sum_j d^2 < max_i sum_{j!=i} d^2`` (Utils.py:169-204)."""
    return _craft_search(own, rows, "std", "sum", layout, metric, search, center, gram_fn, exclude_first)


def opt_fang_attack(own, rows, layout=None, metric="flat", search="reference", center=None,
                    gram_fn=None, exclude_first=True) -> SearchResult:
    """``mu - gamma*sign(mu)`` with the Min-Max acceptance rule (Utils.py:101-132)."""
    return _craft_search(own, rows, "sign", "max", layout, metric, search, center, gram_fn, exclude_first)


def craft(mode: str, own: torch.Tensor, rows: Optional[torch.Tensor], args, *, layout=None,
          valid=None, metric="flat", search="reference", center=None, gram_fn=None,
          strict_reference=False, generator=None) -> Tuple[torch.Tensor, dict]:
    """Dispatch equal to ``RpcClient.malicious_training`` (src/RpcClient.py:119-145)."""
    if mode == "Random":
        if not args:
            raise TypeError("Random attack needs --attack_args <perturbation>")
        return random_attack(own, args[0], valid, generator), {"mode": mode}
    if rows is None or rows.shape[0] == 0:
        return own.clone(), {"mode": mode, "skipped": "no genuine models"}
    if mode == "LIE":
        if not args:
            raise TypeError("LIE attack needs --attack_args <scaling_factor>")
        return lie_attack(rows, args[0], strict_reference), {"mode": mode, "z": float(args[0])}
    fn = {"Min-Max": min_max_attack, "Min-Sum": min_sum_attack, "Opt-Fang": opt_fang_attack}.get(mode)
    if fn is None:
        raise ValueError(f"Attack client not contain '{mode}' algorithm.")
    res = fn(own, rows, layout=layout, metric=metric, search=search, center=center, gram_fn=gram_fn)
    return res.crafted, {"mode": mode, "gamma": res.gamma, "gamma_succ": res.gamma_succ,
                         "threshold": res.threshold, "probes": res.probes}
