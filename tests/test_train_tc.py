"""Tensor-core trainer (``afl_tc_train_round``) vs the FFMA interpreter and a plain PyTorch fp32 reference.

The tcgen05 path multiplies TF32 operands with fp32 accumulation, so gradients agree with fp32 to ~1e-3 relative (max-norm
per tensor); both trainers draw identical Philox dropout masks, so the comparison holds with dropout on."""
import pytest
import torch
import torch.nn.functional as F

from attackfl_b200 import FlatLayout, models
from attackfl_b200.data import make_round_indices, synthetic_icu
from attackfl_b200.models.programs import build_program
from attackfl_b200.ops import train as T
from attackfl_b200.ops import train_tc as TC

pytestmark = pytest.mark.gpu
NAME = "TransformerModel"


def _setup(seed=0):
    torch.manual_seed(seed)
    m = models.build(NAME)
    return m, FlatLayout.from_module(m)


def _interp(lay, rows, ds, idx, n_data, batch, epochs, dropout, **kw):
    S, G = 8, 1
    prog = build_program(NAME, lay, S=S, train=True, dropout=dropout).to("cuda")
    C = rows.shape[0]
    w = rows.clone().cuda()
    m, v = torch.zeros_like(w), torch.zeros_like(w)
    plan = T.TrainPlan(G, (batch + S - 1) // S, S, chunk=C)
    gpart = torch.zeros(C * G, lay.padded, device="cuda")
    loss, steps = T.train_round(prog, plan, w, m, v, gpart, ds.packed().cuda(), idx.cuda(),
                                torch.tensor(n_data, dtype=torch.int32, device="cuda"),
                                torch.arange(C, dtype=torch.int32, device="cuda"), batch, epochs, dropout=dropout, **kw)
    return w.cpu(), loss.cpu(), steps.cpu()


def _tc(lay, rows, ds, idx, n_data, batch, epochs, dropout, **kw):
    C = rows.shape[0]
    w = rows.clone().cuda()
    m, v = torch.zeros_like(w), torch.zeros_like(w)
    gw = torch.zeros(C, lay.padded, device="cuda")
    stash = torch.zeros(C, TC.stash_floats_per_client("cuda"), device="cuda")
    loss, steps = TC.train_round_tc(lay, w, m, v, gw, stash, ds.packed().cuda(), idx.cuda(),
                                    torch.tensor(n_data, dtype=torch.int32, device="cuda"),
                                    torch.arange(C, dtype=torch.int32, device="cuda"), batch, epochs, dropout=dropout, **kw)
    torch.cuda.synchronize()
    return w.cpu(), loss.cpu(), steps.cpu(), gw.cpu()


def _per_tensor_rel(lay, a, b):
    worst = {}
    for e in lay.entries:
        x, y = a[e.offset:e.offset + e.numel], b[e.offset:e.offset + e.numel]
        den = float(y.abs().max())
        if den > 0:
            worst[e.name] = float((x - y).abs().max()) / den
        else:
            assert float(x.abs().max()) == 0.0, e.name
    return worst


@pytest.mark.parametrize("dropout", [False, True])
@pytest.mark.parametrize("batch,n", [(128, 128), (128, 100), (64, 64)])
def test_one_sgd_step_is_the_gradient(dropout, batch, n):
    """lr = 1 SGD without momentum: w0 - w1 is the gradient.  Compared per tensor against the interpreter (same masks)."""
    _, lay = _setup(1)
    assert TC.supported(NAME, lay, batch)
    ds = synthetic_icu(400, 5)
    n_data = [n, n - 3]
    idx = make_round_indices(400, n_data, 1, torch.Generator().manual_seed(3))
    rows = torch.stack([lay.flatten(models.build(NAME).state_dict()) for _ in n_data])
    kw = dict(optimizer="sgd", lr=1.0, momentum=0.0, seed=1234)
    ref, rloss, rsteps = _interp(lay, rows, ds, idx, n_data, batch, 1, dropout, **kw)
    got, loss, steps, gw = _tc(lay, rows, ds, idx, n_data, batch, 1, dropout, **kw)
    assert steps.tolist() == rsteps.tolist()
    for c in range(len(n_data)):
        assert loss[c, 1].item() == 0
        assert loss[c, 0].item() == pytest.approx(rloss[c, 0].item(), rel=2e-3)
        g_ref, g_got = rows[c] - ref[c], rows[c] - got[c]
        worst = _per_tensor_rel(lay, g_got, g_ref)
        bad = {k: v for k, v in worst.items() if v > 1e-2}
        assert not bad, bad
        # the Q / K thirds of in_proj get exactly zero gradient (one token: softmax == 1)
        e = lay.by_name["vitals_transformer.attention.in_proj_weight"]
        assert float(g_got[e.offset:e.offset + 2 * 64 * 64].abs().max()) == 0.0


def test_one_step_matches_autograd_without_dropout():
    m, lay = _setup(2)
    m.eval()
    ds = synthetic_icu(300, 7)
    n_data = [128]
    idx = make_round_indices(300, n_data, 1, torch.Generator().manual_seed(5))
    rows = lay.flatten(m.state_dict())[None]
    got, loss, _, _ = _tc(lay, rows, ds, idx, n_data, 128, 1, False, optimizer="sgd", lr=1.0, momentum=0.0)
    ii = idx[0, 0, :128].long()
    l = F.binary_cross_entropy(m(ds.vitals[ii], ds.labs[ii]), ds.labels[ii][:, None])
    m.zero_grad()
    l.backward()
    ref = lay.flatten({k: (p.grad if p.grad is not None else torch.zeros_like(p)) for k, p in m.named_parameters()})
    assert loss[0, 0].item() / 128 == pytest.approx(l.item(), rel=2e-3)
    worst = _per_tensor_rel(lay, rows[0] - got[0], ref)
    bad = {k: v for k, v in worst.items() if v > 1e-2}
    assert not bad, bad


@pytest.mark.parametrize("dropout", [False, True])
def test_ragged_adam_round_tracks_the_interpreter(dropout):
    """Production shape: batch 128, several epochs, ragged clients (last batches of 44 and of 1, which is skipped)."""
    _, lay = _setup(3)
    ds = synthetic_icu(1500, 11)
    n_data = [300, 257, 1024, 128]
    epochs = 2
    idx = make_round_indices(1500, n_data, epochs, torch.Generator().manual_seed(7))
    rows = torch.stack([lay.flatten(models.build(NAME).state_dict()) for _ in n_data])
    kw = dict(lr=0.004, seed=77)
    ref, rloss, rsteps = _interp(lay, rows, ds, idx, n_data, 128, epochs, dropout, **kw)
    got, loss, steps, _ = _tc(lay, rows, ds, idx, n_data, 128, epochs, dropout, **kw)
    assert steps.tolist() == rsteps.tolist()
    for c in range(len(n_data)):
        assert loss[c, 1].item() == 0
        assert loss[c, 0].item() == pytest.approx(rloss[c, 0].item(), rel=5e-3)
        d_ref, d_got = ref[c] - rows[c], got[c] - rows[c]
        rel = float((d_got - d_ref).norm() / d_ref.norm())
        assert rel < 0.05, (c, rel)


def test_proper_clip_and_sgd_momentum():
    _, lay = _setup(4)
    ds = synthetic_icu(600, 13)
    n_data = [256, 200]
    idx = make_round_indices(600, n_data, 2, torch.Generator().manual_seed(9))
    rows = torch.stack([lay.flatten(models.build(NAME).state_dict()) for _ in n_data])
    kw = dict(optimizer="sgd", lr=0.05, momentum=0.5, clip_proper=True, clip=0.05)
    ref, _, _ = _interp(lay, rows, ds, idx, n_data, 128, 2, False, **kw)
    got, _, _, _ = _tc(lay, rows, ds, idx, n_data, 128, 2, False, **kw)
    for c in range(2):
        d_ref, d_got = ref[c] - rows[c], got[c] - rows[c]
        assert float((d_got - d_ref).norm() / d_ref.norm()) < 1e-2


@pytest.mark.parametrize("cluster", [2, 4, 8])
def test_cluster_shared_optimiser_pass_matches_one_cta(cluster, monkeypatch):
    """The helper CTAs of a client's cluster only take shares of the element-wise optimiser pass, so every cluster size
    must reproduce the one-CTA result up to the run-to-run noise of the kernel's floating-point atomics (column sums and
    the loss are accumulated with shared-memory atomics); a lost or early hand-over would show up as a stale gradient or
    weight, orders of magnitude above that noise.
    Ragged clients (one with a skipped trailing batch of one sample), dropout on, then SGD-momentum with the proper clip
    (the clip coefficient travels to the helpers through distributed shared memory)."""
    _, lay = _setup(5)
    ds = synthetic_icu(1500, 17)
    n_data = [300, 257, 1024, 128, 640]
    epochs = 3
    idx = make_round_indices(1500, n_data, epochs, torch.Generator().manual_seed(11))
    rows = torch.stack([lay.flatten(models.build(NAME).state_dict()) for _ in n_data])
    runs = [dict(lr=0.004, seed=5), dict(optimizer="sgd", lr=0.05, momentum=0.5, clip_proper=True, clip=0.05, seed=6)]
    for kw in runs:
        monkeypatch.setenv("ATTACKFL_TC_CLUSTER", "1")
        ref, rloss, rsteps, _ = _tc(lay, rows, ds, idx, n_data, 128, epochs, True, **kw)
        ref2, _, _, _ = _tc(lay, rows, ds, idx, n_data, 128, epochs, True, **kw)
        assert TC.last_cluster("cuda") == 1
        monkeypatch.setenv("ATTACKFL_TC_CLUSTER", str(cluster))
        got, loss, steps, _ = _tc(lay, rows, ds, idx, n_data, 128, epochs, True, **kw)
        assert TC.last_cluster("cuda") == cluster
        assert steps.tolist() == rsteps.tolist()
        assert torch.allclose(loss, rloss, rtol=1e-4)
        for c in range(len(n_data)):
            upd = float((ref[c] - rows[c]).norm())
            noise = float((ref2[c] - ref[c]).norm()) / upd
            diff = float((got[c] - ref[c]).norm()) / upd
            assert diff <= max(4 * noise, 2e-4), (c, diff, noise)


def test_cluster_size_is_chosen_from_the_client_count(monkeypatch):
    """Few clients: clusters of 8; more clients than the GPU has room for clusters of 8: one CTA per client."""
    monkeypatch.delenv("ATTACKFL_TC_CLUSTER", raising=False)
    _, lay = _setup(6)
    ds = synthetic_icu(400, 19)
    for C, want in ((2, 8), (32, 1), (160, 1)):
        n_data = [128] * C
        idx = make_round_indices(400, n_data, 1, torch.Generator().manual_seed(3))
        rows = torch.stack([lay.flatten(models.build(NAME).state_dict())] * C)
        _tc(lay, rows, ds, idx, n_data, 128, 1, False, lr=0.004)
        assert TC.last_cluster("cuda") == want
