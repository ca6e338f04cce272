"""FlatLayout round-trips, .pth key parity, model key/init parity with the reference (SURVEY §4.3 'Unit')."""
import os

import pytest
import torch

from attackfl_b200 import FlatLayout, models
from attackfl_b200.models.hyper import FusedHyperNet

ICU = ["CNNModel", "RNNModel", "TransformerModel"]


@pytest.mark.parametrize("name,D", [("CNNModel", 203649), ("RNNModel", 97665), ("TransformerModel", 47693)])
def test_param_counts(name, D):
    m = models.build(name)
    lay = FlatLayout.from_module(m)
    assert lay.numel == D
    assert lay.padded % 1024 == 0
    assert all(e.offset % 32 == 0 for e in lay.entries)


@pytest.mark.parametrize("name", ICU + ["TransformerClassifier"])
def test_keys_and_init_match_reference(ref, name):
    torch.manual_seed(3)
    a = models.build(name).state_dict()
    torch.manual_seed(3)
    b = getattr(ref.model, name)().state_dict()
    assert list(a) == list(b)
    for k in a:
        assert torch.equal(a[k], b[k]), k


@pytest.mark.parametrize("name", ICU)
def test_flatten_roundtrip_and_pth(tmp_path, ref, name):
    torch.manual_seed(0)
    m = models.build(name)
    lay = FlatLayout.from_module(m)
    row = lay.flatten(m.state_dict())
    assert row.shape == (lay.padded,)
    assert float(row[~lay.valid_mask()].abs().sum()) == 0.0
    sd = lay.unflatten(row)
    for k, v in m.state_dict().items():
        assert torch.equal(sd[k], v)
    dense = lay.to_dense(row)
    ref_vec = ref.utils.state_dict_to_vector(m.state_dict())
    assert torch.equal(dense, ref_vec)
    assert torch.equal(lay.from_dense(dense), row)
    p = str(tmp_path / f"{name}.pth")
    lay.save_pth(row, p)
    # the reference's module must load our file
    rm = getattr(ref.model, name)()
    rm.load_state_dict(torch.load(p, weights_only=True))
    assert torch.equal(lay.load_pth(p), row)


def test_forward_matches_reference(ref):
    torch.manual_seed(1)
    v, l = torch.randn(9, 7), torch.randn(9, 16)
    v[0, 2] = -2.0
    for name in ICU:
        torch.manual_seed(5)
        a = models.build(name).eval()
        torch.manual_seed(5)
        b = getattr(ref.model, name)().eval()
        assert torch.allclose(a(v, l), b(v, l), atol=1e-7), name
    torch.manual_seed(5)
    a = models.build("TransformerClassifier").eval()
    torch.manual_seed(5)
    b = ref.model.TransformerClassifier().eval()
    x = torch.randn(2, 1, 561)
    assert torch.allclose(a(x), b(x), atol=1e-6)


def test_hypernetwork_parity_and_fused_export(ref):
    torch.manual_seed(0)
    h = models.HyperNetwork(models.TransformerModel(), 3, 8, 100, False, 2)
    torch.manual_seed(0)
    hr = ref.model.HyperNetwork(ref.model.TransformerModel(), 3, 8, 100, False, 2)
    sa, sb = h.state_dict(), hr.state_dict()
    assert list(sa) == list(sb) and len(sa) == 121
    f = FusedHyperNet.from_module(h)
    ex = f.export_state_dict()
    assert list(ex) == list(sb)
    assert all(torch.equal(ex[k], sb[k]) for k in ex)
    w, emd = hr(torch.tensor([2]))
    th = f.emit_torch(torch.tensor([2]))[0]
    got = f.layout.views(th)
    for k in w:
        assert torch.allclose(got[k], w[k], atol=1e-6), k
    assert torch.equal(f.embedding(2), emd)
    # round trip back into a reference-structured module
    h2 = f.to_module(models.TransformerModel())
    assert all(torch.equal(h2.state_dict()[k], sb[k]) for k in sb)


def test_cnnhyper_keys(ref):
    torch.manual_seed(0)
    a = models.CNNHyper(3, 10, 100, 3)
    torch.manual_seed(0)
    b = ref.model.CNNHyper(3, 10, 100, 3)
    assert list(a.state_dict()) == list(b.state_dict())
    wa, _ = a(torch.tensor([1]))
    wb, _ = b(torch.tensor([1]))
    assert all(torch.allclose(wa[k], wb[k]) for k in wb)
