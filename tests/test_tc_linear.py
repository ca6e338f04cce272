"""tcgen05 linear kernels (csrc/gemm_tcgen05.cu) against plain fp32 PyTorch; TF32 operands => ~1e-3 relative."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [(64, 64), (192, 64), (256, 64), (64, 256), (32, 32), (128, 96)]     # (N, K): out_proj, in_proj, ffn1, ffn2


def _close(got, want, tol=4e-3):
    scale = want.abs().max().item() + 1e-12
    err = (got - want).abs().max().item() / scale
    assert err < tol, f"relative error {err:.3e}"


def _mk(m, n, k, seed=0):
    g = torch.Generator(device="cuda").manual_seed(seed)
    x = torch.randn(m, k, device="cuda", generator=g)
    w = torch.randn(n, k, device="cuda", generator=g) / k ** 0.5
    b = torch.randn(n, device="cuda", generator=g)
    dy = torch.randn(m, n, device="cuda", generator=g)
    return x, w, b, dy


@pytest.mark.parametrize("n,k", SHAPES)
@pytest.mark.parametrize("m", [128, 1000, 5 * 561, 300 * 128 + 37])   # last: more tiles than SMs (multi-tile CTAs)
def test_linear_fwd(m, n, k):
    from attackfl_b200.ops import tc_linear as tc
    x, w, b, _ = _mk(m, n, k)
    _close(tc.linear_fwd(x, w, b), x @ w.t() + b)
    _close(tc.linear_fwd(x, w, None, relu=True), torch.relu(x @ w.t()))


@pytest.mark.parametrize("n,k", SHAPES)
@pytest.mark.parametrize("m", [128, 1000, 5 * 561, 300 * 128 + 37])   # last: more tiles than SMs (multi-tile CTAs)
def test_linear_dx_dw(m, n, k):
    from attackfl_b200.ops import tc_linear as tc
    x, w, b, dy = _mk(m, n, k, seed=1)
    _close(tc.linear_dx(dy, w), dy @ w)
    mask = torch.randn(m, k, device="cuda")
    _close(tc.linear_dx(dy, w, mask_src=mask, mask_scale=1.25), (dy @ w) * (mask > 0) * 1.25)
    acc = torch.randn(m, k, device="cuda")
    want = acc + dy @ w
    _close(tc.linear_dx(dy, w, out=acc, accumulate=True), want)
    dw, db = tc.linear_dw(dy, x)
    _close(dw, dy.t() @ x)
    _close(db, dy.sum(0), tol=1e-4)
    dw2, db2 = tc.linear_dw(dy, x, dw=dw.clone(), db=db.clone(), accumulate=True)
    _close(dw2, 2 * (dy.t() @ x))
    _close(db2, 2 * dy.sum(0), tol=1e-4)


def test_linear_dropout_matches_philox_stream():
    from attackfl_b200.ops import tc_linear as tc
    from attackfl_b200.ops import train as nt
    m, n, k = 300, 64, 64
    x, w, b, _ = _mk(m, n, k, seed=2)
    y = tc.linear_fwd(x, w, b, relu=True, drop_p=0.25, seed=99, rng=(3, 7, 11))
    ref = torch.relu(x @ w.t() + b)
    keep = torch.tensor([[float(nt.dropout_keep(99, 3, 7, r, 11, c, 0.25)) for c in range(n)] for r in range(4)],
                        device="cuda")
    _close(y[:4], ref[:4] * keep / 0.75)
    frac = (y != 0).float().sum() / (ref != 0).float().sum()
    assert abs(frac.item() - 0.75) < 0.02


def test_autograd_function_matches_linear():
    from attackfl_b200.ops.tc_linear import TcLinear
    x, w, b, dy = _mk(640, 256, 64, seed=3)
    xs = [t.clone().requires_grad_(True) for t in (x, w, b)]
    y = TcLinear.apply(xs[0], xs[1], xs[2], True)
    y.backward(dy)
    _close(y, torch.relu(torch.nn.functional.linear(x, w, b)))
    g = dy * (y > 0)              # same ReLU mask: pre-activations within TF32 error of 0 may flip in the fp32 oracle
    _close(xs[0].grad, g @ w)
    _close(xs[1].grad, g.t() @ x)
    _close(xs[2].grad, g.sum(0), tol=1e-4)


@pytest.mark.parametrize("k,p", [(64, 0.0), (64, 0.2), (256, 0.1)])
@pytest.mark.parametrize("m", [128, 777, 300 * 128 + 37])
def test_linear_layernorm_fused_epilogue(m, k, p):
    """GEMM + bias + dropout + residual + LayerNorm in one kernel against the unfused kernels (same Philox stream) and,
    without dropout, against torch."""
    import ctypes
    from attackfl_b200.ops import tc_linear as tc
    from attackfl_b200.ops.native import check, cuda_lib, ptr
    x, w, b, _ = _mk(m, 64, k, seed=7)
    g = torch.Generator(device="cuda").manual_seed(3)
    resid, gamma, beta = (torch.randn(s, device="cuda", generator=g) for s in ((m, 64), (64,), (64,)))
    y, xh, rs = torch.empty(m, 64, device="cuda"), torch.empty(m, 64, device="cuda"), torch.empty(m, device="cuda")
    tc.linear_ln_fwd(x, w, b, resid, gamma, beta, y, xh, rs, drop_p=p, seed=5, rng=(2, 9, 18))
    lib = cuda_lib()
    lin = tc.linear_fwd(x, w, b)
    y2, xh2, rs2 = torch.empty_like(y), torch.empty_like(xh), torch.empty_like(rs)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(lib, lib.afl_har_add_ln_fwd(ptr(resid), ptr(lin), ctypes.c_float(p), ctypes.c_uint64(5), 2, 9, 18, ptr(gamma), ptr(beta),
                                      ctypes.c_int64(m), ptr(y2), ptr(xh2), ptr(rs2), st))
    assert torch.allclose(y, y2, atol=2e-5, rtol=1e-5) and torch.allclose(xh, xh2, atol=2e-5, rtol=1e-5)
    assert torch.allclose(rs, rs2, rtol=1e-5)
    if p == 0.0:
        want = torch.nn.functional.layer_norm(resid + x @ w.t() + b, (64,), gamma, beta)
        _close(y, want)
