"""Native kernels vs plain PyTorch fp32 references.  Each test runs on the CPU twin (same kernel
sources compiled with g++) and, marked ``gpu``, on the sm_100a library."""
import math

import pytest
import torch

from attackfl_b200 import FlatLayout, models
from attackfl_b200 import aggregators as agg
from attackfl_b200 import attacks as atk
from attackfl_b200.models.hyper import FusedHyperNet
from attackfl_b200.ops import flat

DEVICES = ["cpu", pytest.param("cuda", marks=pytest.mark.gpu)]


def _rand_rows(c, d, device, seed=0, spread=0.05):
    g = torch.Generator().manual_seed(seed)
    base = torch.randn(d, generator=g)
    rows = base[None] + spread * torch.randn(c, d, generator=g)
    return rows.to(device).contiguous()


@pytest.mark.parametrize("device", DEVICES)
def test_adam_flat_matches_torch(device):
    torch.manual_seed(0)
    w0 = torch.randn(3, 1024)
    p = torch.nn.Parameter(w0.clone())
    opt = torch.optim.Adam([p], lr=0.004)
    w = w0.clone().to(device)
    m, v = torch.zeros_like(w), torch.zeros_like(w)
    for step in range(1, 6):
        g = torch.randn(3, 1024)
        p.grad = g.clone()
        opt.step()
        flat.adam_flat(w, g.to(device), m, v, 0.004, step)
    assert torch.allclose(w.cpu(), p.data, atol=1e-6, rtol=1e-5)


@pytest.mark.parametrize("device", DEVICES)
def test_sgdm_flat_matches_torch(device):
    torch.manual_seed(0)
    w0 = torch.randn(2048)
    p = torch.nn.Parameter(w0.clone())
    opt = torch.optim.SGD([p], lr=0.01, momentum=0.5)
    w = w0.clone().to(device)
    buf = torch.zeros_like(w)
    for step in range(4):
        g = torch.randn(2048)
        p.grad = g.clone()
        opt.step()
        flat.sgdm_flat(w, g.to(device), buf, 0.01, 0.5, step == 0)
    assert torch.allclose(w.cpu(), p.data, atol=1e-6, rtol=1e-5)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("c,a", [(3, 0), (8, 1), (24, 3), (16, 6)])
def test_fedavg_with_moments(device, c, a):
    d = 4096
    rows = _rand_rows(c, d, device, seed=c)
    sizes = torch.randint(12000, 15001, (c,)).float()
    wts = (sizes / sizes.sum()).to(device)
    bcast = torch.zeros(c, d, device=device)
    member = None
    z = None
    if a:
        g = torch.Generator().manual_seed(1)
        member = (torch.rand(a, c, generator=g) < 0.6).float()
        member[:, 0] = 1.0
        member[:, 1] = 1.0
        z = torch.linspace(0.5, 1.5, a)
    out, mu, sd, lie = flat.fedavg(rows, wts, bcast=bcast, member=member, z=z, ref=rows[0].clone(), want_lie=bool(a))
    want = agg.fedavg(rows.cpu(), sizes.tolist())
    assert torch.allclose(out.cpu(), want, atol=1e-5, rtol=1e-5)
    assert torch.allclose(bcast.cpu(), want[None].expand(c, d), atol=1e-5, rtol=1e-5)
    for q in range(a):
        sub = rows.cpu()[member[q] > 0]
        assert torch.allclose(mu[q].cpu(), sub.mean(0), atol=1e-5, rtol=1e-5)
        assert torch.allclose(sd[q].cpu(), sub.std(0), atol=1e-5, rtol=1e-4)
        assert torch.allclose(lie[q].cpu(), atk.lie_attack(sub, float(z[q])), atol=1e-5, rtol=1e-4)


@pytest.mark.parametrize("device", DEVICES)
def test_fedavg_k1_guard(device):
    rows = _rand_rows(4, 1024, device)
    member = torch.tensor([[1.0, 0, 0, 0]])
    wts = torch.full((4,), 0.25, device=device)
    _, mu, sd, _ = flat.fedavg(rows, wts, member=member, want_mu=True, want_sd=True)
    assert torch.equal(sd[0].cpu(), torch.zeros(1024)) and torch.allclose(mu[0], rows[0])
    _, _, sd, _ = flat.fedavg(rows, wts, member=member, want_sd=True, strict_k1=True)
    assert torch.isnan(sd).all()


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("c", [3, 10, 11, 32, 37, 64, 65, 128, 200, 256])
def test_coord_select(device, c):
    rows = _rand_rows(c, 2048, device, seed=c, spread=1.0)
    assert torch.equal(flat.coord_select(rows, "median").cpu(), agg.median(rows.cpu()))
    k = int(c * 0.1)
    got = flat.coord_select(rows, "trimmed", k).cpu()
    assert torch.allclose(got, agg.trimmed_mean(rows.cpu(), 0.1), atol=1e-5, rtol=1e-5)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("mode", ["Min-Max", "Min-Sum", "Opt-Fang"])
def test_gamma_search_and_craft(device, mode):
    torch.manual_seed(3)
    k, d = 9, 4096
    rows = _rand_rows(k, d, "cpu", seed=5, spread=0.03)
    res = {"Min-Max": atk.min_max_attack, "Min-Sum": atk.min_sum_attack, "Opt-Fang": atk.opt_fang_attack}[mode](
        rows[0], rows, metric="flat", center=rows.mean(0))
    mu, sd = atk.mean_std(rows)
    p = sd if mode != "Opt-Fang" else torch.sign(mu)
    c = rows.mean(0)
    a = torch.cat([rows - c, (mu - c)[None], p[None]])
    G = (a.double() @ a.double().t()).float().to(device)
    out = flat.gamma_search(G, k, "sum" if mode == "Min-Sum" else "max")
    assert out[0].item() == pytest.approx(res.gamma)
    assert out[1].item() == pytest.approx(res.gamma_succ)
    assert out[2].item() == pytest.approx(res.threshold, rel=1e-4)
    crafted = flat.craft_axpy(mu.to(device), p.to(device), out, sign_mode=(mode == "Opt-Fang"))
    assert torch.allclose(crafted.cpu(), res.crafted, atol=1e-5)


@pytest.mark.parametrize("device", DEVICES)
def test_craft_random(device):
    w = torch.zeros(1 << 16, device=device)
    valid = torch.ones_like(w)
    valid[-100:] = 0
    out = flat.craft_random(w, valid, 0.5, seed=7, stream_id=3).cpu()
    assert abs(out[:-100].std().item() - 0.5) < 0.01 and abs(out.mean().item()) < 0.01
    assert float(out[-100:].abs().sum()) == 0.0
    again = flat.craft_random(w, valid, 0.5, seed=7, stream_id=3).cpu()
    other = flat.craft_random(w, valid, 0.5, seed=7, stream_id=4).cpu()
    assert torch.equal(out, again) and not torch.equal(out, other)


@pytest.mark.parametrize("device", DEVICES)
def test_hyper_emit_and_step_match_torch(device):
    torch.manual_seed(0)
    n = 4
    h = models.HyperNetwork(models.TransformerModel(), n, 8, 100, False, 2)
    ref = FusedHyperNet.from_module(h, device="cpu")
    fused = FusedHyperNet.from_module(h, device=device)
    nat = flat.NativeHyper(fused)
    lay = fused.layout
    theta = nat.emit(list(range(n)))
    want = ref.emit_torch(torch.arange(n))
    assert torch.allclose(theta.cpu(), want, atol=1e-5)
    rows = []
    for i in range(n):
        torch.manual_seed(20 + i)
        rows.append(lay.flatten(models.TransformerModel().state_dict()))
    rows = torch.stack(rows)
    agg.hyper_round_torch(ref, list(range(n)), rows, 1e-3, 1.0)
    rows_d = rows.to(device)
    for i in range(n):
        nat.step(i, rows_d[i], 1e-3, 1.0)
    nat.unpack()
    got, exp = fused.export_state_dict(), ref.export_state_dict()
    for k in exp:
        assert torch.allclose(got[k], exp[k], atol=3e-6, rtol=1e-4), k
    assert fused.step == ref.step == n


@pytest.mark.parametrize("device", DEVICES)
def test_scion_quantize_and_scores(device):
    """ScionFL pieces (src/Utils.py:372-387, server.py:443-466): the packed-bit pass against the plain fp32 maths
    evaluated on the very same bits, plus the Bernoulli statistics of the quantiser."""
    c, d_pad = 9, 4096
    rows = _rand_rows(c, d_pad, device, seed=4, spread=0.5)
    rows[3] *= 25.0                                            # one client far above the 3x mean L2 cap
    valid = torch.ones(d_pad, device=device)
    valid[-70:] = 0
    rows[:, -70:] = 0
    cosd, aux = flat.scion_cosine_distances(rows, valid, seed=11)
    bits, counts, minmax = aux["bits"].cpu(), aux["counts"].cpu(), aux["minmax"].cpu()
    dense = rows.cpu()[:, :-70]
    assert torch.equal(minmax[:, 0], dense.min(1).values) and torch.equal(minmax[:, 1], dense.max(1).values)
    sigma = ((bits[:, :, None] >> torch.arange(32)) & 1).reshape(c, d_pad).float()
    assert float(sigma[:, -70:].sum()) == 0.0
    sigma = sigma[:, :-70]
    assert torch.equal(sigma.sum(1).int(), counts)
    probs = (dense - minmax[:, :1]) / (minmax[:, 1:] - minmax[:, :1] + 1e-6)
    dev = (sigma.sum(1) - probs.sum(1)).abs() / (probs * (1 - probs)).sum(1).sqrt()
    assert float(dev.max()) < 4.5                              # counts within 4.5 sigma of their expectation
    assert abs(float(((sigma - probs) * (probs - 0.5)).mean())) < 0.01
    lohi = aux["lohi"].cpu().float()
    want = agg.scion_dequant_scores(sigma, lohi[:, 0], lohi[:, 1])
    assert torch.allclose(cosd.cpu(), want, atol=2e-5)
    assert lohi[3, 1] < minmax[3, 1] * 0.9                     # the outlier was clipped
    again, _ = flat.scion_cosine_distances(rows, valid, seed=11)
    other, _ = flat.scion_cosine_distances(rows, valid, seed=12)
    assert torch.allclose(cosd, again, atol=1e-6) and not torch.allclose(cosd, other, atol=1e-7)
    out, kept = agg.scionfl(rows, [1.0] * c, valid=valid.bool(), scores=cosd.tolist())
    assert 1 <= len(kept) < c and out.shape == (d_pad,)


@pytest.mark.parametrize("device", DEVICES)
def test_scion_slice_owned_pieces_reproduce_the_one_pass(device):
    """The N-GPU decomposition of the ScionFL scoring pass: every 'rank' quantises only its own rows (Philox stream id =
    position of the row in the reporting list) and scores only its slice of the packed words, reading every client's
    codes by address; scalars and partial sums added up must give the one-GPU distances."""
    c, d_pad, world = 10, 4096, 2
    rows = _rand_rows(c, d_pad, device, seed=5, spread=0.5)
    valid = torch.ones(d_pad, device=device)
    valid[-40:] = 0
    rows[:, -40:] = 0
    want, aux = flat.scion_cosine_distances(rows, valid, seed=21)
    cpg, words = c // world, d_pad // 32
    bits, counts, minmax = [], [], []
    for r in range(world):                                   # owner-computes: rows r * cpg .. (r + 1) * cpg
        b, n, mm = flat.scion_quantize(rows[r * cpg:(r + 1) * cpg].contiguous(), valid, 21, row0=r * cpg)
        bits.append(b.contiguous()); counts.append(n); minmax.append(mm)
    assert torch.equal(torch.cat(bits), aux["bits"]) and torch.equal(torch.cat(counts), aux["counts"])
    counts, minmax = torch.cat(counts), torch.cat(minmax)
    d = float(valid.sum().item())
    lo, hi = flat.scion_clipped_levels(counts, minmax, d)
    lohi = torch.stack([lo, hi], 1)
    ptrs = [bits[i // cpg][i % cpg].data_ptr() for i in range(c)]
    sums = sum(flat.scion_scores_ptrs(ptrs, words, r * words // world, (r + 1) * words // world, valid, lohi)
               for r in range(world))
    got = flat.scion_distances_from_sums(sums, counts, lo, hi, d)
    assert torch.allclose(got, want, atol=2e-6)


@pytest.mark.parametrize("device", DEVICES)
def test_round_indices_are_keyed_permutations(device):
    """``afl_round_indices``: every (client, epoch) row is a permutation of ONE subset of the pool (sampling without
    replacement, a fresh order per epoch), reproducible from (seed, client id), different across clients / seeds, and
    statistically uniform: every record is picked with probability size / n_rows and lands anywhere in the epoch."""
    n_rows, epochs = 1000, 3
    sizes = torch.tensor([400, 257, 1, 1000, 640], dtype=torch.int32, device=device)
    cids = torch.tensor([0, 1, 2, 9, 4], dtype=torch.int32, device=device)
    idx = flat.round_indices(n_rows, sizes, cids, epochs, 1000, seed=77).cpu()
    assert idx.shape == (5, epochs, 1000) and idx.dtype == torch.int32
    for c, n in enumerate(sizes.tolist()):
        first = set(idx[c, 0, :n].tolist())
        assert len(first) == n and min(first) >= 0 and max(first) < n_rows
        for e in range(epochs):
            assert (idx[c, e, n:] == -1).all()
            assert set(idx[c, e, :n].tolist()) == first                      # the same subset every epoch ...
        if n > 8:
            assert idx[c, 0, :n].tolist() != idx[c, 1, :n].tolist()           # ... in a different order
    again = flat.round_indices(n_rows, sizes, cids, epochs, 1000, seed=77).cpu()
    other = flat.round_indices(n_rows, sizes, cids, epochs, 1000, seed=78).cpu()
    assert torch.equal(idx, again) and not torch.equal(idx, other)
    # the CPU twin and the GPU evaluate the same integer function
    ref = flat.round_indices(n_rows, sizes.cpu(), cids.cpu(), epochs, 1000, seed=77)
    assert torch.equal(idx, ref)
    # statistics over 400 clients: inclusion frequency of every record ~ Binomial(400, 0.3); mean position of a record
    # inside the epochs it appears in ~ uniform
    C = 400
    big = flat.round_indices(n_rows, torch.full((C,), 300, dtype=torch.int32, device=device),
                             torch.arange(C, dtype=torch.int32, device=device), 2, 300, seed=5).cpu().long()
    counts = torch.bincount(big[:, 0, :].reshape(-1), minlength=n_rows).float()
    z = (counts - C * 0.3) / (C * 0.3 * 0.7) ** 0.5
    assert float(z.abs().max()) < 5.0 and abs(float(z.mean())) < 0.2 and 0.8 < float(z.std()) < 1.2
    pos = torch.arange(300).repeat(C).float()
    mean_pos = torch.zeros(n_rows).index_add_(0, big[:, 1, :].reshape(-1), pos) / counts.clamp_min(1)
    assert abs(float(mean_pos[counts > 0].mean()) - 149.5) < 2.0
    # consecutive positions are not correlated with consecutive records
    d = (big[:, 0, 1:] - big[:, 0, :-1]).float()
    assert abs(float(d.mean())) < 3.0 and float((d == 1).float().mean()) < 0.01


@pytest.mark.parametrize("device", DEVICES)
def test_native_roc_auc_matches_the_torch_restatement(device):
    """``afl_roc_auc`` (one block: bitonic sort of 64-bit keys, prefix counts, tie-group scan, trapezoid) against
    ``validation.roc_auc`` (== sklearn ``auc(roc_curve)``), with heavy ties, tiny sets, a single class and NaN scores."""
    from attackfl_b200.validation import roc_auc
    g = torch.Generator().manual_seed(3)
    for n in (2, 3, 31, 100, 1000, 1024, 1025, 4096, 5000, 8192):
        s = torch.rand(n, generator=g)
        s[::3] = (s[::3] * 8).round() / 8                      # large tie groups, also across classes
        y = (torch.rand(n, generator=g) < 0.3).float()
        y[0], y[1] = 1.0, 0.0
        out = flat.roc_auc(s.to(device), y.to(device)).cpu()
        assert float(out[0]) == pytest.approx(roc_auc(s, y), abs=2e-6), n
        assert out[1] == 0 and out[2] == y.sum() and out[3] == n - y.sum()
    const = flat.roc_auc(torch.full((64,), 0.5, device=device), (torch.arange(64) % 2).float().to(device)).cpu()
    assert float(const[0]) == pytest.approx(0.5)               # all scores tied: chance level
    perfect = flat.roc_auc(torch.arange(10, device=device).float(), (torch.arange(10) >= 5).float().to(device)).cpu()
    assert float(perfect[0]) == pytest.approx(1.0)
    one_class = flat.roc_auc(torch.rand(10).to(device), torch.ones(10, device=device)).cpu()
    assert one_class[0] != one_class[0] and one_class[1] == 0
    bad = torch.rand(50)
    bad[7] = float("nan")
    out = flat.roc_auc(bad.to(device), torch.ones(50, device=device)).cpu()
    assert out[0] != out[0] and out[1] == 1
