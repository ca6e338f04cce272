"""One Gram for all attackers (``ops.flat.multi_gram_attack``, csrc/gram_multi.cu) vs the per-attacker Gram path."""
import pytest
import torch

from attackfl_b200.ops import flat, gram

pytestmark = pytest.mark.gpu


def _one(sub, cen, mode, prec):
    kk = sub.shape[0]
    _, mu, sd, _ = flat.fedavg(sub.contiguous(), torch.zeros(kk, device="cuda"), member=torch.ones(1, kk), ref=cen.contiguous(),
                               want_mu=True, want_sd=True)
    p = torch.sign(mu[0]) if mode == "Opt-Fang" else sd[0]
    a = torch.cat([sub - cen, (mu[0] - cen)[None], p[None]], dim=0)
    G = gram.gram(a, precision=prec)
    g4 = flat.gamma_search(G, kk, "sum" if mode == "Min-Sum" else "max", True, False)
    return flat.craft_axpy(mu[0].contiguous(), p.contiguous(), g4), g4


@pytest.mark.parametrize("n,D,prec", [(12, 4096, "tf32"), (48, 98304, "tf32"), (112, 48128, "fp32")])
def test_all_attackers_from_one_gram(n, D, prec):
    torch.manual_seed(n)
    cen = torch.randn(D, device="cuda") * 0.3
    rows = cen[None] + 0.02 * torch.randn(n, D, device="cuda")
    modes = ["Min-Max", "Min-Sum", "Opt-Fang", "Min-Max", "Opt-Fang", "Min-Sum", "Min-Max"]
    g = torch.Generator().manual_seed(1)
    k = max(n // 2, 2)
    subs = [torch.randperm(n, generator=g)[:k].tolist() for _ in modes]
    out, gam = flat.multi_gram_attack([rows[j].data_ptr() for j in range(n)], cen, subs, modes, D, precision=prec)
    torch.cuda.synchronize()
    for a, mode in enumerate(modes):
        want, g4 = _one(rows[subs[a]], cen, mode, prec)
        assert float(gam[a, 0]) == float(g4[0]), (a, mode, gam[a].tolist(), g4.tolist())       # same bisection outcome
        assert float(gam[a, 2]) == pytest.approx(float(g4[2]), rel=2e-3)
        assert torch.allclose(out[a], want, atol=1e-5, rtol=1e-5), (a, mode)


def test_sliced_calls_cover_the_vector():
    """Two coordinate slices with the partial Grams added = the unsliced call (what two ranks do)."""
    torch.manual_seed(0)
    n, D = 10, 8192
    cen = torch.randn(D, device="cuda") * 0.1
    rows = cen[None] + 0.05 * torch.randn(n, D, device="cuda")
    modes, subs = ["Min-Max", "Min-Sum"], [[0, 3, 5, 7, 9], [1, 2, 3, 4, 8]]
    ptrs = [rows[j].data_ptr() for j in range(n)]
    full, gam = flat.multi_gram_attack(ptrs, cen, subs, modes, D, precision="fp32")
    parts = []

    def collect(G):
        parts.append(G.clone())
        return G
    for lo, hi in ((0, D // 2), (D // 2, D)):
        flat.multi_gram_attack(ptrs, cen, subs, modes, D, lo=lo, hi=hi, precision="fp32", reduce_gram=collect)
    Gsum = parts[0] + parts[1]
    out = torch.zeros(2, D, device="cuda")
    for lo, hi in ((0, D // 2), (D // 2, D)):
        dst = [[out[a].data_ptr()] for a in range(2)]
        flat.multi_gram_attack(ptrs, cen, subs, modes, D, lo=lo, hi=hi, precision="fp32", dst_ptrs=dst,
                               reduce_gram=lambda G: Gsum)
    torch.cuda.synchronize()
    assert torch.allclose(out, full, atol=1e-5, rtol=1e-5)
