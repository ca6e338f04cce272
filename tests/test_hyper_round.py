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
