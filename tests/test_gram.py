"""K3: tcgen05/TMEM/TMA split-K Gram kernel vs plain PyTorch fp32 (GPU only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rows(r, d, seed=0, spread=0.05):
    g = torch.Generator().manual_seed(seed)
    base = torch.randn(d, generator=g)
    return (base[None] * 0 + spread * torch.randn(r, d, generator=g)).cuda().contiguous()


@pytest.mark.parametrize("r,d", [(5, 4096), (18, 48128), (50, 98304), (128, 98304), (130, 48128), (256, 203776),
                                 (33, 12345)])
def test_gram_matches_fp32_of_bf16_operands(r, d):
    from attackfl_b200.ops import gram_native
    a = _rows(r, d, seed=r)
    g = gram_native.gram_tcgen05(a, "bf16")
    ab = a.to(torch.bfloat16).to(torch.float64)
    want = (ab @ ab.t()).float()
    # same operands (bf16-rounded), fp32 accumulation in TMEM vs fp64: only accumulation-order error remains
    assert torch.allclose(g, want, rtol=2e-4, atol=2e-4 * float(want.abs().max())), (g - want).abs().max()


@pytest.mark.parametrize("r,d", [(18, 48128), (50, 98304)])
def test_gram_distances_close_to_fp32(r, d):
    """Centred bf16 operands: pairwise distances within 1e-3 relative of the fp32 reference."""
    from attackfl_b200.ops import gram_native
    a = _rows(r, d, seed=3)
    g = gram_native.gram_tcgen05(a, "bf16").double()
    dg = torch.diagonal(g)
    d_bf = (dg[:, None] + dg[None, :] - 2 * g).clamp_min(0).sqrt()
    d_ref = torch.cdist(a.double(), a.double())
    mask = ~torch.eye(r, dtype=torch.bool, device="cuda")
    rel = ((d_bf - d_ref).abs() / d_ref.clamp_min(1e-12))[mask]
    assert float(rel.max()) < 1e-3
    g3 = gram_native.gram_tcgen05(a, "bf16x3").double()
    d3 = torch.diagonal(g3)
    d_3 = (d3[:, None] + d3[None, :] - 2 * g3).clamp_min(0).sqrt()
    assert float(((d_3 - d_ref).abs() / d_ref.clamp_min(1e-12))[mask].max()) < 2e-5


@pytest.mark.parametrize("r,d", [(5, 4096), (18, 48128), (50, 98304), (128, 98304), (130, 48128), (256, 203776),
                                 (33, 12348)])
def test_gram_tf32_direct(r, d):
    """kind::tf32 path: fp32 rows straight from the caller's matrix (TMA zero-fills rows >= R and the K tail)."""
    from attackfl_b200.ops import gram_native
    a = _rows(r, d, seed=r + 1)
    g = gram_native.gram_tcgen05(a, "tf32")
    at = (a.view(torch.int32) & -8192).view(torch.float32).double()       # tf32 = fp32 with 13 mantissa bits dropped
    want = (at @ at.t()).float()
    exact = (a.double() @ a.double().t()).float()
    scale = float(exact.abs().max())
    # the tensor core reads the top 19 bits of each fp32 operand (truncation) or rounds: accept either
    err_t, err_e = float((g - want).abs().max()), float((g - exact).abs().max())
    assert min(err_t, err_e) < 2e-4 * scale and err_e < 3e-3 * scale, (err_t, err_e, scale)
    gd = g.double()
    dg = torch.diagonal(gd)
    dist = (dg[:, None] + dg[None, :] - 2 * gd).clamp_min(0).sqrt()
    ref = torch.cdist(a.double(), a.double())
    mask = ~torch.eye(r, dtype=torch.bool, device="cuda")
    assert float(((dist - ref).abs() / ref.clamp_min(1e-12))[mask].max()) < 1e-3
