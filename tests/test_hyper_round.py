"""Persistent hyper-network round (``afl_hyper_round`` / ``afl_hyper_emit_rows``, csrc/hyper_round.cu) vs the torch oracle
(``aggregators.hyper_round_torch``, itself checked against the reference's ``Server.train_hyper``)."""
import copy

import pytest
import torch

from attackfl_b200 import FlatLayout, models
from attackfl_b200.aggregators import hyper as H
from attackfl_b200.models.hyper import FusedHyperNet, HyperNetwork
from attackfl_b200.ops.flat import NativeHyper

pytestmark = pytest.mark.gpu


def _net(n_clients, model="TransformerModel", seed=0):
    torch.manual_seed(seed)
    target = models.build(model)
    lay = FlatLayout.from_module(target)
    hnet = HyperNetwork(target, n_clients, 8, 100, False, 2)
    return lay, FusedHyperNet.from_module(hnet, lay, device="cuda")


@pytest.mark.parametrize("model,n", [("TransformerModel", 5), ("RNNModel", 3)])
def test_round_matches_sequential_torch_steps(model, n):
    lay, fused = _net(n, model)
    ref = copy.deepcopy(fused)
    torch.manual_seed(1)
    rows = torch.stack([lay.flatten(models.build(model).state_dict()) for _ in range(n)]).cuda()
    order = [2, 0, 1] if n == 3 else [0, 1, 2, 3, 4]
    for rnd in range(2):                                  # two rounds: Adam step counter and moments carry over
        nat = NativeHyper(fused) if rnd == 0 else nat
        norms = nat.round(order, [rows[c].data_ptr() for c in order], 1e-3, 1.0)
        want = H.hyper_round_torch(ref, order, rows[order], 1e-3, 1.0, "sequential")
        torch.cuda.synchronize()
        assert nat.check_round_error() == 0
        nat.unpack()
        assert torch.allclose(norms.cpu(), torch.tensor(want), rtol=1e-4)
        for k, v in ref.trainable().items():
            got = fused.trainable()[k]
            assert torch.allclose(got, v, atol=2e-6, rtol=1e-4), (rnd, k, float((got - v).abs().max()))
        for k in ref.m:
            assert torch.allclose(fused.m[k], ref.m[k], atol=1e-7, rtol=1e-3), k
        assert fused.step == ref.step
        rows = rows + 0.01 * torch.randn_like(rows)


def test_sharded_rows_give_the_same_update_as_one_shard():
    """Two launches on disjoint row shards (what two ranks do, minus the cross-rank sum) cannot be compared directly because
    the scalars are shard-local; instead one shard covering everything must equal the unsharded call, and emit_rows must
    equal the dense emit on any shard."""
    lay, fused = _net(4)
    nat = NativeHyper(fused)
    ids = [3, 1, 0]
    dense = nat.emit(ids)
    out = torch.zeros(len(ids), lay.padded, device="cuda")
    half = lay.padded // 2 // 64 * 64
    for lo, hi in ((0, half), (half, lay.padded)):
        nat.setup_round(lo, hi)
        nat.emit_rows(ids, [out[j].data_ptr() for j in range(len(ids))])
    torch.cuda.synchronize()
    assert torch.allclose(out, dense, atol=1e-6)


def test_two_emulated_ranks_share_one_round(monkeypatch):
    """Row-sharded round on ONE GPU: two "ranks" (disjoint row shards, concurrent cooperative launches on two streams) exchange
    their H + 1 partial sums through the flag protocol, exactly as two GPUs do (minus the NVLink hop).  The result must equal
    the unsharded round."""
    import ctypes
    monkeypatch.setenv("ATTACKFL_FUSED_MAX_BLOCKS", "60")        # both grids must be co-resident
    n = 4
    lay, fused_a = _net(n)
    fused_b, fused_ref = copy.deepcopy(fused_a), copy.deepcopy(fused_a)
    torch.manual_seed(3)
    rows = torch.stack([lay.flatten(models.build("TransformerModel").state_dict()) for _ in range(n)]).cuda()
    order = [1, 3, 0, 2]
    ptrs = [rows[c].data_ptr() for c in order]
    ref = NativeHyper(fused_ref)
    ref.round(order, ptrs, 1e-3, 1.0)
    half = lay.padded // 2
    H = fused_a.H
    xch = [torch.zeros(n * 2 * (H + 1), device="cuda") for _ in range(2)]
    sig = [torch.zeros(n * 2, dtype=torch.int32, device="cuda") for _ in range(2)]
    PA = ctypes.c_void_p * 2
    xp, sp = PA(*[t.data_ptr() for t in xch]), PA(*[t.data_ptr() for t in sig])
    nats = [NativeHyper(fused_a), NativeHyper(fused_b)]
    nats[0].setup_round(0, half, 2, 0, n, xp, sp)
    nats[1].setup_round(half, lay.padded, 2, 1, n, xp, sp)
    streams = [torch.cuda.Stream(), torch.cuda.Stream()]
    torch.cuda.synchronize()
    for r in range(2):
        with torch.cuda.stream(streams[r]):
            nats[r].round(order, ptrs, 1e-3, 1.0)
    torch.cuda.synchronize()
    for r in range(2):
        assert nats[r].check_round_error() == 0
        nats[r].unpack()
    ref.unpack()
    lo_hi = ((0, half), (half, lay.padded))
    for r, f in enumerate((fused_a, fused_b)):
        lo, hi = lo_hi[r]
        assert torch.allclose(f.head_w[lo:hi], fused_ref.head_w[lo:hi], atol=2e-6, rtol=1e-4)
        assert torch.allclose(f.head_b[lo:hi], fused_ref.head_b[lo:hi], atol=2e-6, rtol=1e-4)
        for k in ("emb", "mlp_w0", "mlp_w2", "mlp_b1"):          # the replicated small parameters stay identical on both ranks
            assert torch.allclose(f.trainable()[k], fused_ref.trainable()[k], atol=2e-6, rtol=1e-4), k
