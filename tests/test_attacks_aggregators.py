"""Algorithm parity vs the reference's ``src.Utils`` / ``server.py`` maths (SURVEY §4.3 'Algorithm parity')."""
import copy
import math

import numpy as np
import pytest
import torch

from attackfl_b200 import FlatLayout, models
from attackfl_b200 import aggregators as agg
from attackfl_b200 import attacks as atk


def _rows(name, k, spread=0.05, seed=0):
    torch.manual_seed(seed)
    base = models.build(name)
    lay = FlatLayout.from_module(base)
    w = lay.flatten(base.state_dict())
    rows = w[None] + spread * torch.randn(k, lay.padded) * lay.valid_mask()[None]
    return lay, w, rows


def _sds(lay, rows):
    return [lay.unflatten(r) for r in rows]


@pytest.mark.parametrize("k", [2, 5, 12])
def test_lie_matches_reference(ref, k):
    lay, w, rows = _rows("RNNModel", k)
    got = atk.lie_attack(rows, 0.74)
    want = lay.flatten(ref.utils.create_LIE_state_dict(_sds(lay, rows), 0.74))
    assert torch.allclose(got, want, atol=2e-6, rtol=1e-5)


def test_lie_k1_guard_and_strict():
    lay, w, rows = _rows("TransformerModel", 1)
    assert torch.isfinite(atk.lie_attack(rows, 0.74)).all()
    assert torch.isnan(atk.lie_attack(rows, 0.74, strict_reference=True)).any()


@pytest.mark.parametrize("mode,fn", [("Min-Max", "create_min_max_model"), ("Min-Sum", "create_min_sum_model"),
                                     ("Opt-Fang", "create_opt_fang_model")])
@pytest.mark.parametrize("name", ["TransformerModel", "RNNModel"])
def test_distance_attacks_reference_metric(ref, mode, fn, name):
    """metric='reference' must reproduce the reference bit-for-bit in gamma and ~1e-6 in weights."""
    lay, w, rows = _rows(name, 4, spread=0.02, seed=2)
    own = lay.unflatten(w)
    want = lay.flatten(getattr(ref.utils, fn)(own, _sds(lay, rows)))
    got, info = atk.craft(mode, w, rows, None, layout=lay, metric="reference")
    assert torch.allclose(got, want, atol=1e-5, rtol=1e-5)
    assert info["gamma"] == pytest.approx(1.5625) or info["gamma_succ"] > 0


@pytest.mark.parametrize("mode", ["Min-Max", "Min-Sum", "Opt-Fang"])
def test_flat_metric_gram_equals_explicit(mode):
    """The Gram formulation equals brute-force flat-L2 distances (fp64 oracle)."""
    lay, w, rows = _rows("TransformerModel", 7, spread=0.03, seed=4)
    res = {"Min-Max": atk.min_max_attack, "Min-Sum": atk.min_sum_attack, "Opt-Fang": atk.opt_fang_attack}[mode](
        w, rows, layout=lay, metric="flat", center=w)
    r64 = rows.double()
    mu = r64.mean(0)
    p = r64.std(0) if mode != "Opt-Fang" else torch.sign(mu)
    pair = torch.cdist(r64, r64)
    thr = pair.max().item() if mode != "Min-Sum" else (pair ** 2).sum(1).max().item()
    assert res.threshold == pytest.approx(thr, rel=1e-4)
    for gamma, ok in res.probes:
        m = mu - gamma * p
        d = (r64[1:] - m).norm(dim=1)
        want = (d.max().item() < thr) if mode != "Min-Sum" else ((d ** 2).sum().item() < thr)
        assert ok == want
    assert torch.allclose(res.crafted.double(), mu - res.gamma * p, atol=1e-5)


def test_cnn_works_in_flat_and_layerfro_but_not_reference():
    lay, w, rows = _rows("CNNModel", 3, spread=0.01)
    for metric in ("flat", "layer-fro"):
        out, info = atk.craft("Min-Max", w, rows, None, layout=lay, metric=metric)
        assert torch.isfinite(out).all()
    with pytest.raises(RuntimeError):
        atk.craft("Min-Max", w, rows, None, layout=lay, metric="reference")


def test_single_genuine_model_returns_own():
    lay, w, rows = _rows("TransformerModel", 1)
    out, _ = atk.craft("Min-Sum", w, rows, None, layout=lay)
    assert torch.equal(out, w)


def test_gamma_schedule_paper_vs_reference():
    g, succ, probes = atk.gamma_schedule(lambda x: x < 7.0)
    assert [p[0] for p in probes] == [50.0, 25.0, 12.5, 6.25, 9.375, 7.8125]
    assert g == 7.8125 and succ == 6.25
    g2, succ2, _ = atk.gamma_schedule(lambda x: x < 7.0, search="paper")
    assert g2 == succ2 == 6.25


def test_random_attack_statistics():
    lay, w, _ = _rows("TransformerModel", 1)
    g = torch.Generator().manual_seed(0)
    out = atk.random_attack(w, 0.5, lay.valid_mask(), g)
    d = lay.to_dense(out - w)
    assert abs(d.std().item() - 0.5) < 0.02
    assert float((out * ~lay.valid_mask()).abs().sum()) == 0.0


# ------------------------------------------------------------------ aggregators
def test_fedavg_matches_reference_math(ref):
    lay, w, rows = _rows("RNNModel", 6)
    sizes = [12000, 15000, 13000, 12500, 14999, 12001]
    sds = _sds(lay, rows)
    want = {k: sum(sds[i][k] * sizes[i] for i in range(6)) / sum(sizes) for k in sds[0]}
    assert torch.allclose(agg.fedavg(rows, sizes), lay.flatten(want), atol=1e-6)


def test_trimmed_mean_and_median_match_reference(ref):
    lay, w, rows = _rows("TransformerModel", 11)
    sds = _sds(lay, rows)
    assert torch.allclose(agg.trimmed_mean(rows, 0.1), lay.flatten(ref.utils.trimmed_mean_aggregation(copy.deepcopy(sds))), atol=1e-6)
    assert torch.equal(agg.median(rows), lay.flatten(ref.utils.median_aggregation(sds)))
    lay, w, rows = _rows("TransformerModel", 10)   # even n => lower median
    assert torch.equal(agg.median(rows), lay.flatten(ref.utils.median_aggregation(_sds(lay, rows))))


def test_krum_matches_reference(ref):
    lay, w, rows = _rows("TransformerModel", 9, spread=0.05)
    rows[3] += 0.5
    dense = [lay.to_dense(r) for r in rows]
    want = ref.utils.krum(dense, 0)
    i = agg.krum_select(rows, 0)
    assert torch.equal(dense[i], want)


def test_shieldfl_matches_reference_formula():
    lay, w, rows = _rows("TransformerModel", 6, spread=0.05)
    vecs = [r / (r.norm() + 1e-8) for r in rows]
    refv = sum(vecs) / len(vecs)
    cos = torch.tensor([torch.nn.functional.cosine_similarity(v[None], refv[None]).item() for v in vecs])
    wts = 1 / (1 - cos + 1e-6)
    wts = wts / wts.sum()
    # 1/(1-cos) amplifies fp32 rounding of cos ~ 1 - 1e-4; compare loosely on weights, tightly on result
    got = agg.shieldfl_weights(rows)
    assert torch.allclose(got, wts, rtol=5e-2)
    assert torch.allclose(agg.shieldfl(rows), (rows * wts[:, None]).sum(0), atol=1e-3)


def test_scionfl_runs_and_keeps_half():
    lay, w, rows = _rows("TransformerModel", 8, spread=0.05)
    out, kept = agg.scionfl(rows, [100] * 8, valid=lay.valid_mask(), generator=torch.Generator().manual_seed(0))
    assert 1 <= len(kept) <= 8 and torch.isfinite(out).all()


def test_fltrust_downweights_opposite_update():
    lay, w, rows = _rows("TransformerModel", 5, spread=0.0)
    d = 0.01 * torch.randn(lay.padded) * lay.valid_mask()
    rows = w[None] + d[None] * torch.tensor([1.0, 1.1, 0.9, 1.0, -1.0])[:, None]
    new, ts = agg.fltrust(rows, w, w + d)
    assert ts[4] == 0 and ts[:4].min() > 0.99
    assert torch.allclose(new, w + d, atol=1e-5)


def test_p2_filters():
    lay, w, rows = _rows("TransformerModel", 10, spread=0.01)
    rows[7] += 1.0
    assert 7 in agg.fltracer_anomalies(rows)
    kept = agg.gmm_filter(rows, [False] * 10)
    assert len(kept) >= 5
    assert torch.isfinite(agg.byzantine_tolerance(rows)).all()


def test_hyper_step_matches_reference_train_hyper(ref):
    """One sequential round of the fused closed-form == autograd.grad + clip + Adam of server.py:637-678."""
    from collections import OrderedDict
    from attackfl_b200.models.hyper import FusedHyperNet
    torch.manual_seed(0)
    n = 3
    hr = ref.model.HyperNetwork(ref.model.TransformerModel(), n, 8, 100, False, 2)
    fused = FusedHyperNet.from_module(hr)
    lay = fused.layout
    opt = torch.optim.Adam(hr.parameters(), lr=1e-3)
    client_rows = []
    for i in range(n):
        torch.manual_seed(10 + i)
        client_rows.append(lay.flatten(ref.model.TransformerModel().state_dict()))
    for node in range(n):
        weights, _ = hr(torch.tensor([node]))
        inner = OrderedDict({k: t.data for k, t in weights.items()})
        opt.zero_grad()
        final = lay.views(client_rows[node])
        delta = OrderedDict({k: inner[k] - final[k] for k in weights})
        grads = torch.autograd.grad(list(weights.values()), hr.parameters(), grad_outputs=list(delta.values()),
                                    allow_unused=True)
        for p, g in zip(hr.parameters(), grads):
            p.grad = g
        torch.nn.utils.clip_grad_norm_(hr.parameters(), 1.0)
        opt.step()
    agg.hyper_round_torch(fused, list(range(n)), torch.stack(client_rows), 1e-3, 1.0)
    want = hr.state_dict()
    got = fused.export_state_dict()
    for k in want:
        assert torch.allclose(got[k], want[k], atol=2e-6, rtol=1e-5), k


def test_krum_vectorised_matches_the_reference_loop():
    """Vectorised top-k Krum against the reference's Python loop (src/Utils.py:326-342), f = 0 and f > 0."""
    import numpy as np
    torch.manual_seed(3)
    from attackfl_b200.aggregators.core import krum_select
    for n, f in ((7, 0), (9, 2), (3, 0), (2, 0)):
        rows = torch.randn(n, 50)
        vecs = [r.numpy() for r in rows]
        scores = []
        for i in range(n):
            d = sorted(float(np.linalg.norm(vecs[i] - vecs[j]) ** 2) for j in range(n) if j != i)
            scores.append(sum(d[:n - f - 2]))
        assert krum_select(rows, f) == int(np.argmin(scores))
    # negative n - f - 2: the reference's slice keeps len(d) - |n-f-2| distances, the default clamps to 0
    rows = torch.randn(5, 20)
    vecs = [r.numpy() for r in rows]
    f = 4
    scores = [sum(sorted(float(np.linalg.norm(vecs[i] - vecs[j]) ** 2) for j in range(5) if j != i)[:5 - f - 2]) for i in range(5)]
    assert krum_select(rows, f, reference_slice=True) == int(np.argmin(scores))
    assert krum_select(rows, f) == 0


def test_gram_falls_back_above_256_rows():
    from attackfl_b200.ops.gram import gram
    a = torch.randn(300, 64)
    assert torch.allclose(gram(a), a @ a.t(), atol=1e-4)
