"""Server-side aggregators / defences on flat rows ``[n, D_pad]``.

Reference parity (all in ``/root/reference``): FedAvg ``server.py:751-775``; trimmed mean
``src/Utils.py:267-302``; median ``:344-357``; Krum ``:326-342`` + ``server.py:373-389``; ShieldFL
``server.py:306-350``; ScionFL ``server.py:436-492`` + ``src/Utils.py:372-387``; FLTrust
``server.py:703-743``; GMM ``server.py:352-372`` + ``src/Utils.py:257-262,304-323`` (broken in the
reference); FLTracer ``server.py:395-435`` (commented out) + ``src/Utils.py:363-369``;
``byzantine_tolerance_aggregation`` ``src/Utils.py:228-248`` (never called).

These are the pure-torch forms (oracle, CPU path, small-n fall-back).  The engine swaps in the
native kernels where they exist: K1 (weighted reduce+scale+broadcast), K3 Gram for every pairwise
/ cosine quantity, K6 for per-coordinate selection; see ``attackfl_b200/ops``.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, Tuple

import numpy as np
import torch

GramFn = Callable[[torch.Tensor], torch.Tensor]


def _gram(rows: torch.Tensor, gram_fn: Optional[GramFn] = None) -> torch.Tensor:
    if gram_fn is not None:
        return gram_fn(rows)
    r = rows.to(torch.float32)
    return r @ r.t()


# ---------------------------------------------------------------------------- FedAvg
def fedavg(rows: torch.Tensor, sizes: Sequence[float]) -> torch.Tensor:
    """``sum_i n_i w_i / sum_i n_i`` (server.py:766-769)."""
    w = torch.as_tensor(sizes, dtype=torch.float32, device=rows.device)
    return (rows * w[:, None]).sum(0) / w.sum()


def weighted_sum(rows: torch.Tensor, weights: torch.Tensor) -> torch.Tensor:
    return (rows * weights.to(rows.dtype).to(rows.device)[:, None]).sum(0)


# ---------------------------------------------------------------------------- coordinate-wise
def trimmed_mean(rows: torch.Tensor, trim_ratio: float = 0.1) -> torch.Tensor:
    n = rows.shape[0]
    k = int(n * trim_ratio)
    if 2 * k >= n:
        raise ValueError("too few clients for the chosen trim_ratio")
    s, _ = torch.sort(rows, dim=0)
    return s[k:n - k].mean(0)


def median(rows: torch.Tensor) -> torch.Tensor:
    """Lower median for even n (``torch.median`` semantics, Utils.py:355-356)."""
    return torch.median(rows, dim=0).values


# ---------------------------------------------------------------------------- Gram-based
def pairwise_sqdist(rows: torch.Tensor, gram_fn: Optional[GramFn] = None,
                    center: bool = True) -> torch.Tensor:
    r = rows - rows.mean(0, keepdim=True) if center else rows
    g = _gram(r, gram_fn).to(torch.float64)
    d = torch.diagonal(g)
    return (d[:, None] + d[None, :] - 2 * g).clamp_min(0)


def krum_select(rows: torch.Tensor, f: int = 0, gram_fn: Optional[GramFn] = None, reference_slice: bool = False) -> int:
    """Index of the Krum winner: argmin_i of the sum of the ``n-f-2`` smallest ``||v_i-v_j||^2`` (one top-k, one argmin;
    no per-client host sync).

    ``reference_slice``: keep the reference's slice semantics for ``n-f-2 < 0`` (``sorted(d)[:n-f-2]`` with a negative
    bound keeps the first ``len(d) - |n-f-2|`` distances, ``src/Utils.py:326-342``); otherwise the count is clamped at 0.
    """
    n = rows.shape[0]
    d2 = pairwise_sqdist(rows, gram_fn)
    m = n - f - 2
    if m < 0:
        m = max((n - 1) + m, 0) if reference_slice else 0
    m = min(m, n - 1)
    if m == 0 or n < 2:
        return 0                                            # every score is 0: argmin returns the first client
    off = d2.clone()
    off.fill_diagonal_(float("inf"))                        # a client's distance to itself is not a candidate
    scores = torch.topk(off, m, dim=1, largest=False).values.sum(1)
    return int(torch.argmin(scores))


def shieldfl_weights(rows: torch.Tensor, gram_fn: Optional[GramFn] = None) -> torch.Tensor:
    """ShieldFL aggregation weights from the Gram of the normalised rows:
    ``<v_i, ref> = mean_j G_ij``, ``||ref||^2 = mean_ij G``; ``a_i ~ 1/(1-cos_i+1e-6)``."""
    g = _gram(rows, gram_fn).to(torch.float64)
    nrm = torch.diagonal(g).clamp_min(0).sqrt() + 1e-8
    gh = g / (nrm[:, None] * nrm[None, :])
    vi_ref = gh.mean(1)
    ref_norm = gh.mean().clamp_min(0).sqrt()
    vi_norm = torch.diagonal(gh).clamp_min(0).sqrt()
    cos = vi_ref / torch.clamp(vi_norm * ref_norm, min=1e-8)
    a = 1.0 / (1.0 - cos + 1e-6)
    return (a / a.sum()).to(torch.float32)


def shieldfl(rows: torch.Tensor, gram_fn: Optional[GramFn] = None) -> torch.Tensor:
    return weighted_sum(rows, shieldfl_weights(rows, gram_fn))


def fltrust(rows: torch.Tensor, global_row: torch.Tensor, server_row: torch.Tensor,
            gram_fn: Optional[GramFn] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Published FLTrust (Cao et al.): trust = ReLU(cos(delta_i, delta_0)),
    ``w <- w + sum_i TS_i * (||d0||/(||d_i||+1e-6)) d_i / (sum TS + 1e-6)``.

    The reference's implementation aliases ``g_0`` to live parameters and subtracts the global
    model twice, which makes its update a no-op; the formula and epsilons are the
    reference's (server.py:714-742), the aliasing bugs are not reproduced.
    """
    deltas = torch.cat([rows - global_row, (server_row - global_row)[None]], dim=0)
    g = _gram(deltas, gram_fn).to(torch.float64)
    n = rows.shape[0]
    norms = torch.diagonal(g).clamp_min(0).sqrt()
    cos = g[:n, n] / torch.clamp(norms[:n] * norms[n], min=1e-8)
    ts = cos.clamp_min(0)
    scale = (norms[n] / (norms[:n] + 1e-6)) * ts
    upd = weighted_sum(deltas[:n], scale.to(torch.float32)) / (ts.sum().item() + 1e-6)
    return global_row + upd, ts.to(torch.float32)


# ---------------------------------------------------------------------------- ScionFL
def scion_quantize(vec: torch.Tensor, generator: Optional[torch.Generator] = None):
    """1-bit stochastic quantisation (Utils.py:372-376)."""
    smin, smax = vec.min(), vec.max()
    probs = (vec - smin) / (smax - smin + 1e-6)
    sigma = torch.bernoulli(probs, generator=generator)
    return sigma, smin.item(), smax.item()


def scionfl(rows: torch.Tensor, sizes: Sequence[float], valid: Optional[torch.Tensor] = None,
            generator: Optional[torch.Generator] = None, keep: str = "reference",
            mu_threshold: float = 3.0, topk_ratio: float = 0.5,
            scores: Optional[Sequence[float]] = None) -> Tuple[torch.Tensor, List[int]]:
    """ScionFL-style filter then size-weighted FedAvg of the kept clients (server.py:436-492).

    ``keep='reference'`` keeps clients whose cosine *distance* to the aggregate is above the
    median, i.e. the most dissimilar half (likely inverted in the reference);
    ``keep='similar'`` keeps the closest half.  ``scores``: cosine distances already computed by the
    bit-packed native pass (``ops.flat.scion_cosine_distances``); otherwise the plain torch restatement below.
    """
    if scores is None:
        scores = scion_scores_torch(rows, valid, generator, mu_threshold)
    kept = scionfl_keep(scores, keep, topk_ratio)
    sz = torch.as_tensor([sizes[i] for i in kept], dtype=torch.float32, device=rows.device)
    out = (rows[kept] * sz[:, None]).sum(0) / sz.sum()
    return out, kept


def scionfl_keep(scores: Sequence[float], keep: str = "reference", topk_ratio: float = 0.5) -> List[int]:
    """The keep rule of server.py:467-480 on the cosine distances (see ``scionfl``)."""
    scores = [float(x) for x in scores]
    n = len(scores)
    thr = sorted(scores, reverse=True)[int(topk_ratio * n)]
    if keep == "reference":
        kept = [i for i, s in enumerate(scores) if s > thr]
    else:
        kept = [i for i, s in enumerate(scores) if s <= thr]
    return kept or list(range(n))


def scion_dequant_scores(sigma: torch.Tensor, lo: torch.Tensor, hi: torch.Tensor) -> torch.Tensor:
    """Cosine distances of dequantised rows ``lo + sigma (hi - lo)`` to their mean (server.py:455-466)."""
    deq = lo[:, None] + sigma * (hi - lo)[:, None]
    agg = deq.mean(0)
    return 1 - torch.nn.functional.cosine_similarity(deq, agg[None], dim=1)


def scion_scores_torch(rows: torch.Tensor, valid: Optional[torch.Tensor] = None,
                       generator: Optional[torch.Generator] = None, mu_threshold: float = 3.0) -> List[float]:
    n = rows.shape[0]
    dense = rows if valid is None else rows[:, valid]
    quant = [scion_quantize(dense[i], generator) for i in range(n)]
    d = dense.shape[1]
    l2 = []
    for sigma, smin, smax in quant:
        c1 = sigma.sum().item()
        l2.append(float(np.sqrt((d - c1) * smin ** 2 + c1 * smax ** 2)))
    l2_avg = float(np.mean(l2))
    lo, hi = [], []
    for (sigma, smin, smax), l in zip(quant, l2):
        if l > mu_threshold * l2_avg:
            fct = (mu_threshold * l2_avg) / l
            smin, smax = smin * fct, smax * fct
        lo.append(smin); hi.append(smax)
    sig = torch.stack([q[0] for q in quant])
    return scion_dequant_scores(sig, torch.tensor(lo, device=rows.device), torch.tensor(hi, device=rows.device)).tolist()


# ---------------------------------------------------------------------------- P2: gmm / fltracer / byzantine
def _pca_project(rows: torch.Tensor, q: int) -> np.ndarray:
    x = (rows - rows.mean(0, keepdim=True)).to(torch.float64).cpu()
    g = x @ x.t()
    evals, evecs = torch.linalg.eigh(g)
    idx = torch.argsort(evals, descending=True)[:q]
    return (evecs[:, idx] * evals[idx].clamp_min(0).sqrt()).numpy()


def gmm_filter(rows: torch.Tensor, attacker_mask: Sequence[bool], n_components: int = 2,
               q: int = 4) -> List[int]:
    """GMM + Mahalanobis 3-sigma filter (server.py:352-372) made workable: the reference fits a
    full covariance in D~5e4 dimensions (18 GB) and crashes on an empty malicious list; we fit the
    same 2-component full-covariance mixture on the top-``q`` principal coordinates (computed from
    the n x n Gram matrix) and keep the points within 3 sigma of the benign Mahalanobis spread."""
    from sklearn.mixture import GaussianMixture
    n = rows.shape[0]
    q = max(1, min(q, n - 1))
    z = _pca_project(rows, q)
    if n <= n_components:
        return list(range(n))
    gm = GaussianMixture(n_components=n_components, covariance_type="full", reg_covar=1e-6,
                         random_state=0).fit(z)
    benign = [i for i in range(n) if not attacker_mask[i]] or list(range(n))

    def md(i, c):
        dlt = z[i] - gm.means_[c]
        return float(np.sqrt(max(dlt @ np.linalg.inv(gm.covariances_[c]) @ dlt, 0.0)))

    thr = 3 * np.std([md(i, 0) for i in benign]) if len(benign) > 1 else np.inf
    kept = []
    for i in range(n):
        c = int(np.argmax(gm.predict_proba(z[i:i + 1])[0]))
        if md(i, c) <= thr or not np.isfinite(thr):
            kept.append(i)
    return kept


def fltracer_anomalies(rows: torch.Tensor, threshold: float = 2.5) -> List[int]:
    """PCA(1) + MAD z-score (Utils.py:363-369), PCA via the Gram matrix."""
    from scipy.stats import median_abs_deviation
    z = _pca_project(rows, 1)[:, 0]
    mad = median_abs_deviation(z)
    med = np.median(z)
    scores = np.abs(z - med) / (1.4826 * mad + 1e-6)
    return [int(i) for i in np.where(scores > threshold)[0]]


def byzantine_tolerance(rows: torch.Tensor, threshold: float = 0.9) -> torch.Tensor:
    """Keep rows with cos >= threshold to row 0, unweighted mean (Utils.py:228-248)."""
    cos = torch.nn.functional.cosine_similarity(rows, rows[0:1], dim=1)
    kept = cos >= threshold
    if not bool(kept.any()):
        kept = torch.ones_like(kept)
    return rows[kept].mean(0)
