"""Persistent training kernel vs PyTorch autograd / torch.optim (fp32 references).

The CPU twin runs the *same* kernel source as the GPU build; GPU variants are marked ``gpu``."""
import pytest
import torch
import torch.nn.functional as F

from attackfl_b200 import FlatLayout, models
from attackfl_b200.data import make_round_indices, synthetic_icu
from attackfl_b200.models.programs import build_program
from attackfl_b200.ops import train as T

DEVICES = ["cpu", pytest.param("cuda", marks=pytest.mark.gpu)]
ICU = ["TransformerModel", "RNNModel", "CNNModel"]


def _setup(name, seed=0):
    torch.manual_seed(seed)
    m = models.build(name)
    lay = FlatLayout.from_module(m)
    return m, lay


@pytest.mark.parametrize("name", ICU)
def test_tile_gradients_match_autograd(name):
    m, lay = _setup(name)
    m.eval()
    ds = synthetic_icu(200, 3)
    data = ds.packed()
    prog = build_program(name, lay, train=True, dropout=False)
    w = lay.flatten(m.state_dict())
    for nv in (prog.S, max(2, prog.S - 3)):
        idx = torch.randperm(200)[:prog.S].to(torch.int32)
        gw, _, loss = T.debug_tile(prog, w, data, idx, nv, 1.0 / nv)
        ii = idx[:nv].long()
        l = F.binary_cross_entropy(m(ds.vitals[ii], ds.labs[ii]), ds.labels[ii][:, None])
        m.zero_grad()
        l.backward()
        ref = lay.flatten({k: (p.grad if p.grad is not None else torch.zeros_like(p)) for k, p in m.named_parameters()})
        assert loss[0].item() / nv == pytest.approx(l.item(), rel=1e-5)
        assert torch.allclose(gw, ref, atol=1e-6, rtol=1e-4)
    # structural zeros of SURVEY §0.8: W_hh / the Q,K thirds of in_proj never receive gradient
    if name == "RNNModel":
        e = lay.by_name["vitals_gru1.weight_hh_l0"]
        assert float(gw[e.offset:e.offset + e.numel].abs().sum()) == 0.0
    if name == "TransformerModel":
        e = lay.by_name["labs_transformer.attention.in_proj_weight"]
        assert float(gw[e.offset:e.offset + 2 * 64 * 64].abs().sum()) == 0.0


def _masks(prog, seed, client, step, S):
    """op-order list of (rng, p, kind, group) for ops that draw dropout."""
    out = []
    for row in prog.ops.tolist():
        p = torch.tensor([row[16]], dtype=torch.int32).view(torch.float32).item()
        if p > 0:
            out.append((row[17], p, row[0], row[15]))
    return sorted(out)          # rng ids follow the branch-sequential build order (vitals, labs, head)


def _mask(seed, client, step, S, rng, p, n):
    m = torch.zeros(S, n)
    for s in range(S):
        for e in range(n):
            if T.dropout_keep(seed, client, step, s, rng, e, p):
                m[s, e] = 1.0 / (1.0 - p)
    return m


def _ln(x, sd, key):
    return F.layer_norm(x, (x.shape[-1],), sd[key + ".weight"], sd[key + ".bias"])


def _forward_with_masks(name, sd, v, l, mk):
    """T=1 closed forms with explicit dropout masks, differentiable w.r.t. ``sd``."""
    it = iter(mk)
    if name == "TransformerModel":
        outs = []
        for b, x in (("vitals", v), ("labs", l)):
            t = f"{b}_transformer"
            h0 = F.gelu(x @ sd[f"{b}_dense.weight"].t() + sd[f"{b}_dense.bias"])
            val = h0 @ sd[f"{t}.attention.in_proj_weight"][128:].t() + sd[f"{t}.attention.in_proj_bias"][128:]
            val = val * next(it).repeat_interleave(16, dim=1)
            att = val @ sd[f"{t}.attention.out_proj.weight"].t() + sd[f"{t}.attention.out_proj.bias"]
            y1 = _ln(h0 + att * next(it), sd, f"{t}.attention_norm")
            f1 = F.gelu(y1 @ sd[f"{t}.ffn.0.weight"].t() + sd[f"{t}.ffn.0.bias"]) * next(it)
            f2 = f1 @ sd[f"{t}.ffn.3.weight"].t() + sd[f"{t}.ffn.3.bias"]
            y2 = _ln(y1 + f2 * next(it), sd, f"{t}.ffn_norm")
            outs.append(_ln(y2, sd, f"{b}_bn"))
        c = torch.cat(outs, 1)
        c = F.gelu(c @ sd["fc1.weight"].t() + sd["fc1.bias"]) * next(it)
        c = F.gelu(c @ sd["fc2.weight"].t() + sd["fc2.bias"])
        return torch.sigmoid(c @ sd["output.weight"].t() + sd["output.bias"])
    if name == "RNNModel":
        outs = []
        for b, x in (("vitals", v), ("labs", l)):
            x = torch.where(x == -2.0, torch.zeros_like(x), x)
            for layer in (1, 2, 3):
                hs = []
                for suf in ("", "_reverse"):
                    g = f"{b}_gru{layer}"
                    gi = x @ sd[f"{g}.weight_ih_l0{suf}"].t() + sd[f"{g}.bias_ih_l0{suf}"]
                    bh = sd[f"{g}.bias_hh_l0{suf}"]
                    r = torch.sigmoid(gi[:, :32] + bh[:32])
                    z = torch.sigmoid(gi[:, 32:64] + bh[32:64])
                    n = torch.tanh(gi[:, 64:] + r * bh[64:])
                    hs.append((1 - z) * n)
                x = torch.cat(hs, 1)
            outs.append(_ln(x, sd, f"{b}_ln") * next(it))
        c = torch.cat(outs, 1)
        c = F.relu(c @ sd["fc1.weight"].t() + sd["fc1.bias"])
        c = F.relu(c @ sd["fc2.weight"].t() + sd["fc2.bias"])
        return torch.sigmoid(c @ sd["output.weight"].t() + sd["output.bias"])
    outs = []
    for b, x in (("vitals", v), ("labs", l)):
        x = x[:, None, :]
        for i in (1, 2, 3):
            x = F.relu(F.conv1d(x, sd[f"{b}_conv{i}.weight"], sd[f"{b}_conv{i}.bias"], padding=1))
        outs.append(F.adaptive_avg_pool1d(x, 4).flatten(1) * next(it))
    c = torch.cat(outs, 1)
    for k in ("fc1", "fc2", "fc3"):
        c = F.relu(c @ sd[f"{k}.weight"].t() + sd[f"{k}.bias"])
    return torch.sigmoid(c @ sd["output.weight"].t() + sd["output.bias"])


@pytest.mark.parametrize("name", ICU)
def test_tile_gradients_with_dropout_masks(name):
    """Forward and backward must use the same Philox masks; compare against autograd with those masks."""
    m, lay = _setup(name, seed=1)
    ds = synthetic_icu(64, 5)
    data = ds.packed()
    prog = build_program(name, lay, S=4, train=True, dropout=True)
    S = prog.S
    w = lay.flatten(m.state_dict())
    seed, client, step = 1234567, 3, 17
    idx = torch.arange(S, dtype=torch.int32)
    gw, _, loss = T.debug_tile(prog, w, data, idx, S, 1.0 / S, seed=seed, client=client, step=step, train=True)
    spec = _masks(prog, seed, client, step, S)
    widths = {"TransformerModel": [4, 64, 6, 64, 4, 64, 6, 64, 64], "RNNModel": [64, 64], "CNNModel": [512, 512]}[name]
    assert len(spec) == len(widths)
    mk = [_mask(seed, client, step, S, rng, p, n) for (rng, p, _, _), n in zip(spec, widths)]
    assert 0.0 < sum(float((x == 0).float().mean()) for x in mk) / len(mk) < 0.5
    sd = {k: v.clone().requires_grad_(True) for k, v in m.state_dict().items()}
    out = _forward_with_masks(name, sd, ds.vitals[:S], ds.labs[:S], mk)
    l = F.binary_cross_entropy(out, ds.labels[:S][:, None])
    l.backward()
    ref = lay.flatten({k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in sd.items()})
    assert loss[0].item() / S == pytest.approx(l.item(), rel=1e-5)
    assert torch.allclose(gw, ref, atol=1e-6, rtol=1e-4)


def _torch_round(name, lay, row, ds, idx, n, batch, epochs, lr, opt="adam", momentum=0.0, clip=0.0):
    m = models.build(name)
    m.load_state_dict(lay.unflatten(row))
    m.eval()                                   # dropout off; no BN in these models
    o = torch.optim.Adam(m.parameters(), lr=lr) if opt == "adam" else torch.optim.SGD(m.parameters(), lr=lr, momentum=momentum)
    tot, steps = 0.0, 0
    for e in range(epochs):
        for s0 in range(0, n, batch):
            ii = idx[e, s0:min(n, s0 + batch)].long()
            o.zero_grad()
            if ii.numel() == 1:
                continue
            l = F.binary_cross_entropy(m(ds.vitals[ii], ds.labs[ii]), ds.labels[ii][:, None])
            l.backward()
            if clip > 0:
                torch.nn.utils.clip_grad_norm_(m.parameters(), clip)
            o.step()
            tot += l.item() * ii.numel()
            steps += 1
    return lay.flatten(m.state_dict()), tot, steps


def _native_round(name, lay, rows, ds, idx, n_data, batch, epochs, device, G, M, S, chain=False, **kw):
    prog = build_program(name, lay, S=S, train=True, dropout=False, chain=chain).to(device)
    C = rows.shape[0]
    w = rows.clone().to(device)
    m, v = torch.zeros_like(w), torch.zeros_like(w)
    plan = T.TrainPlan(G, M, S, chunk=C)
    gpart = torch.zeros(C * G, lay.padded, device=device)
    loss, steps = T.train_round(prog, plan, w, m, v, gpart, ds.packed().to(device), idx.to(device),
                                torch.tensor(n_data, dtype=torch.int32, device=device),
                                torch.arange(C, dtype=torch.int32, device=device), batch, epochs, dropout=False, **kw)
    return w.cpu(), loss.cpu(), steps.cpu()


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("chain", [False, True])
@pytest.mark.parametrize("name,S,G,M,batch", [("TransformerModel", 8, 2, 2, 32), ("TransformerModel", 16, 1, 1, 16),
                                              ("RNNModel", 8, 4, 1, 32), ("CNNModel", 4, 2, 2, 16)])
def test_round_matches_torch_adam(device, name, S, G, M, batch, chain):
    if chain and device == "cuda" and S > 8:
        pytest.skip("chain mode keeps every gradient buffer alive: S=16 does not fit in 227 KB of shared memory")
    m0, lay = _setup(name, seed=2)
    ds = synthetic_icu(300, 9)
    n_data = [batch * 2 + 1, batch * 2 + 5, batch * 3]     # first client ends on a batch of 1 (skipped)
    epochs = 2
    g = torch.Generator().manual_seed(0)
    idx = make_round_indices(300, n_data, epochs, g)
    rows = torch.stack([lay.flatten(models.build(name).state_dict()) for _ in n_data])
    got, loss, steps = _native_round(name, lay, rows, ds, idx, n_data, batch, epochs, device, G, M, S, chain=chain, lr=0.004)
    for c, n in enumerate(n_data):
        want, tot, st = _torch_round(name, lay, rows[c], ds, idx[c], n, batch, epochs, 0.004)
        assert int(steps[c]) == st
        assert loss[c, 1].item() == 0
        assert loss[c, 0].item() == pytest.approx(tot, rel=1e-4)
        # Adam normalises gradients, so fp summation-order noise on ~0 gradients can move a coordinate by up to lr per
        # step: require that virtually all coordinates agree tightly and that none is off by more than two steps
        diff = (got[c] - want).abs()
        assert float((diff > 2e-4).float().mean()) < 2e-3 and float(diff.max()) < 2 * 0.004, (c, diff.max())


@pytest.mark.parametrize("device", DEVICES)
def test_round_sgd_momentum_and_proper_clip(device):
    name = "TransformerModel"
    _, lay = _setup(name, seed=4)
    ds = synthetic_icu(200, 1)
    n_data, batch, epochs = [40, 56], 16, 2
    idx = make_round_indices(200, n_data, epochs, torch.Generator().manual_seed(1))
    rows = torch.stack([lay.flatten(models.build(name).state_dict()) for _ in n_data])
    got, _, _ = _native_round(name, lay, rows, ds, idx, n_data, batch, epochs, device, 2, 1, 8,
                              optimizer="sgd", lr=0.05, momentum=0.5, clip_proper=True, clip=0.05)
    for c, n in enumerate(n_data):
        want, _, _ = _torch_round(name, lay, rows[c], ds, idx[c], n, batch, epochs, 0.05, "sgd", 0.5, clip=0.05)
        assert torch.allclose(got[c], want, atol=2e-4, rtol=1e-3)


@pytest.mark.parametrize("device", DEVICES)
def test_split_invariance_with_dropout(device):
    """Dropout masks are keyed by (client, step, sample-in-batch): results must not depend on G/M/S."""
    name = "RNNModel"
    _, lay = _setup(name, seed=6)
    ds = synthetic_icu(150, 2)
    n_data, batch, epochs = [50, 64], 32, 1
    idx = make_round_indices(150, n_data, epochs, torch.Generator().manual_seed(2))
    rows = torch.stack([lay.flatten(models.build(name).state_dict()) for _ in n_data])
    res = []
    for (S, G, M, chain) in ((16, 2, 1, False), (8, 1, 4, False), (8, 4, 1, False), (8, 4, 1, True)) + (((16, 1, 2, True),) if device == "cpu" else ()):
        prog = build_program(name, lay, S=S, train=True, dropout=True, chain=chain).to(device)
        w = rows.clone().to(device)
        m, v = torch.zeros_like(w), torch.zeros_like(w)
        gpart = torch.zeros(2 * G, lay.padded, device=device)
        T.train_round(prog, T.TrainPlan(G, M, S, 2), w, m, v, gpart, ds.packed().to(device), idx.to(device),
                      torch.tensor(n_data, dtype=torch.int32, device=device),
                      torch.arange(2, dtype=torch.int32, device=device), batch, epochs, seed=99, dropout=True)
        res.append(w.cpu())
    assert not torch.allclose(res[0], rows, atol=1e-4)
    for r in res[1:]:
        assert torch.allclose(res[0], r, atol=1e-5)


@pytest.mark.parametrize("device", DEVICES)
@pytest.mark.parametrize("name", ICU)
def test_infer_matches_module(device, name):
    m, lay = _setup(name, seed=8)
    m.eval()
    ds = synthetic_icu(133, 4)
    prog = build_program(name, lay, S={"TransformerModel": 32, "RNNModel": 16, "CNNModel": 8}[name], train=False).to(device)
    rows = torch.stack([lay.flatten(m.state_dict()), lay.flatten(models.build(name).state_dict())]).to(device)
    probs = T.infer(prog, rows, ds.packed().to(device)).cpu()
    with torch.no_grad():
        want = m(ds.vitals, ds.labs)[:, 0]
    assert probs.shape == (2, 133)
    assert torch.allclose(probs[0], want, atol=1e-5)
    assert not torch.allclose(probs[1], want, atol=1e-3)


@pytest.mark.gpu
def test_production_plan_round_matches_torch_adam():
    """The plan bench.py runs (8 clients, 16 blocks per client, tiles of 8 samples, batch 128) against torch.optim.Adam
    on the PyTorch module, dropout off; then dropout on: the same plan against a different split of the batch (the
    Philox masks are keyed by sample, so the result may not depend on G / M / S)."""
    name, device = "TransformerModel", "cuda"
    _, lay = _setup(name, seed=8)
    ds = synthetic_icu(1400, 21)
    n_data = [300, 257, 384, 129, 512, 640, 200, 1024]       # ragged; clients 1 and 3 end on a batch of one sample (skipped)
    batch, epochs = 128, 2
    idx = make_round_indices(1400, n_data, epochs, torch.Generator().manual_seed(4))
    rows = torch.stack([lay.flatten(models.build(name).state_dict()) for _ in n_data])
    got, loss, steps = _native_round(name, lay, rows, ds, idx, n_data, batch, epochs, device, 16, 1, 8, lr=0.004)
    for c, n in enumerate(n_data):
        want, tot, st = _torch_round(name, lay, rows[c], ds, idx[c], n, batch, epochs, 0.004)
        assert int(steps[c]) == st
        assert loss[c, 1].item() == 0 and loss[c, 0].item() == pytest.approx(tot, rel=1e-4)
        diff = (got[c] - want).abs()
        assert float((diff > 2e-4).float().mean()) < 2e-3 and float(diff.max()) < 2 * 0.004, (c, diff.max())
    res = []
    for (S, G, M) in ((8, 16, 1), (8, 4, 4), (16, 8, 1)):
        prog = build_program(name, lay, S=S, train=True, dropout=True).to(device)
        w = rows.clone().to(device)
        m, v = torch.zeros_like(w), torch.zeros_like(w)
        gpart = torch.zeros(len(n_data) * G, lay.padded, device=device)
        T.train_round(prog, T.TrainPlan(G, M, S, len(n_data)), w, m, v, gpart, ds.packed().to(device), idx.to(device),
                      torch.tensor(n_data, dtype=torch.int32, device=device),
                      torch.arange(len(n_data), dtype=torch.int32, device=device), batch, epochs, seed=5, dropout=True)
        res.append(w.cpu())
    assert not torch.allclose(res[0], got, atol=1e-4)          # dropout did change the trajectory
    for r in res[1:]:
        d = (res[0] - r).abs()
        assert float((d > 2e-4).float().mean()) < 2e-3 and float(d.max()) < 2 * 0.004
