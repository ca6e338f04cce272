"""Native HAR trainer (tcgen05 linears + har_kernels.cu) against the PyTorch module (fp32 autograd)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()


def _setup(B=6, seed=0):
    from attackfl_b200 import FlatLayout, models
    from attackfl_b200.ops.har import HarNativeTrainer
    torch.manual_seed(seed)
    mod = models.TransformerClassifier().cuda()
    for m in mod.modules():                       # dropout off: masks of the two implementations are different streams
        if isinstance(m, torch.nn.Dropout):
            m.p = 0.0
        if isinstance(m, torch.nn.MultiheadAttention):
            m.dropout = 0.0
    mod.train()
    lay = FlatLayout.from_module(mod) if hasattr(FlatLayout, "from_module") else FlatLayout(mod.state_dict())
    row = lay.flatten(mod.state_dict(), device="cuda")
    tr = HarNativeTrainer(lay, "cuda", max_batch=max(B, 8))
    x = torch.randn(B, 561, device="cuda")
    y = torch.randint(0, 6, (B,), device="cuda")
    return mod, lay, row, tr, x, y


def test_attention_kernels_match_sdpa():
    from attackfl_b200.ops import har
    from attackfl_b200.ops.native import cuda_lib, ptr, check
    import ctypes
    B, T = 3, 561
    lib = cuda_lib()
    qkv = torch.randn(B * T, 192, device="cuda")
    out = torch.empty(B * T, 64, device="cuda"); lse = torch.zeros(B * 4 * lib.afl_har_lse_stride(T), device="cuda")
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    check(lib, lib.afl_har_attn_fwd(ptr(qkv), B, T, ctypes.c_float(0.0), ctypes.c_uint64(0), 0, 0, 0, ptr(out), ptr(lse), st))
    q, k, v = (qkv[:, i * 64:(i + 1) * 64].reshape(B, T, 4, 16).transpose(1, 2) for i in range(3))
    qr, kr, vr = (t.detach().clone().requires_grad_(True) for t in (q, k, v))
    ref = torch.nn.functional.scaled_dot_product_attention(qr, kr, vr)
    assert _rel(out.reshape(B, T, 4, 16).transpose(1, 2), ref) < 1e-4
    dout = torch.randn(B * T, 64, device="cuda")
    dqkv = torch.zeros(B * T, 192, device="cuda")
    check(lib, lib.afl_har_attn_bwd(ptr(qkv), ptr(out), ptr(dout), ptr(lse), B, T, ctypes.c_float(0.0), ctypes.c_uint64(0), 0,
                                    0, 0, ptr(dqkv), st))
    ref.backward(dout.reshape(B, T, 4, 16).transpose(1, 2))
    for i, gr in enumerate((qr.grad, kr.grad, vr.grad)):
        got = dqkv[:, i * 64:(i + 1) * 64].reshape(B, T, 4, 16).transpose(1, 2)
        assert _rel(got, gr) < 2e-4, i


@pytest.mark.parametrize("T,B", [(561, 3), (200, 2), (128, 1), (37, 2)])
def test_attention_tensor_core_forward(T, B):
    """tcgen05 attention forward against SDPA (no dropout) and against the FFMA kernel with the same dropout stream."""
    from attackfl_b200.ops.native import cuda_lib, ptr, check
    import ctypes
    lib = cuda_lib()
    assert lib.afl_har_attn_tc_supported(T)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    qkv = torch.randn(B * T, 192, device="cuda")

    def run(fn, p):
        tp = lib.afl_har_lse_stride(T)
        o = torch.full((B * T, 64), float("nan"), device="cuda"); l = torch.zeros(B * 4 * tp, device="cuda")
        if fn is lib.afl_har_attn_fwd_tc:
            check(lib, fn(ptr(qkv), B, T, ctypes.c_float(p), ctypes.c_uint64(9), 2, 3, 4, ptr(o), ptr(l), None, st))
        else:
            check(lib, fn(ptr(qkv), B, T, ctypes.c_float(p), ctypes.c_uint64(9), 2, 3, 4, ptr(o), ptr(l), st))
        return o, l
    o_tc, l_tc = run(lib.afl_har_attn_fwd_tc, 0.0)
    q, k, v = (qkv[:, i * 64:(i + 1) * 64].reshape(B, T, 4, 16).transpose(1, 2) for i in range(3))
    ref = torch.nn.functional.scaled_dot_product_attention(q, k, v)
    assert _rel(o_tc.reshape(B, T, 4, 16).transpose(1, 2), ref) < 4e-3
    o_ff, l_ff = run(lib.afl_har_attn_fwd, 0.0)
    assert (l_tc - l_ff).abs().max().item() < 2e-2
    o_tc, _ = run(lib.afl_har_attn_fwd_tc, 0.2)
    o_ff, _ = run(lib.afl_har_attn_fwd, 0.2)
    assert _rel(o_tc, o_ff) < 6e-3


@pytest.mark.parametrize("T,B,p", [(561, 2, 0.0), (561, 2, 0.2), (200, 1, 0.1), (64, 2, 0.0), (37, 1, 0.3)])
def test_attention_tensor_core_backward(T, B, p):
    """tcgen05 attention backward (dq + dkv kernels, keep bits from the forward's mask words) against the FFMA backward
    kernels, which regenerate the same dropout stream; p = 0 is also checked against SDPA autograd."""
    from attackfl_b200.ops.native import cuda_lib, ptr, check
    import ctypes
    lib = cuda_lib()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    tp, mw = lib.afl_har_lse_stride(T), lib.afl_har_attn_mask_words(T)
    qkv = torch.randn(B * T, 192, device="cuda")
    dout = torch.randn(B * T, 64, device="cuda")
    out = torch.empty(B * T, 64, device="cuda"); lse = torch.zeros(B * 4 * tp, device="cuda")
    mask = torch.zeros(B * 4 * mw * tp, dtype=torch.int32, device="cuda")
    delta = torch.zeros(B * 4 * tp, device="cuda")
    cp = ctypes.c_float(p)
    check(lib, lib.afl_har_attn_fwd_tc(ptr(qkv), B, T, cp, ctypes.c_uint64(9), 2, 3, 4, ptr(out), ptr(lse), ptr(mask), st))
    got = torch.full((B * T, 192), float("nan"), device="cuda")
    check(lib, lib.afl_har_attn_bwd_tc(ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(mask), B, T, cp, ptr(delta), ptr(got), st))
    want = torch.zeros(B * T, 192, device="cuda")
    check(lib, lib.afl_har_attn_bwd(ptr(qkv), ptr(out), ptr(dout), ptr(lse), B, T, cp, ctypes.c_uint64(9), 2, 3, 4, ptr(want), st))
    torch.cuda.synchronize()
    for i, name in enumerate("qkv"):
        assert _rel(got[:, i * 64:(i + 1) * 64], want[:, i * 64:(i + 1) * 64]) < 1.5e-2, name   # TF32 dS, P operands
    if p == 0.0:
        q, k, v = (qkv[:, i * 64:(i + 1) * 64].reshape(B, T, 4, 16).transpose(1, 2).detach().clone().requires_grad_(True)
                   for i in range(3))
        torch.nn.functional.scaled_dot_product_attention(q, k, v).backward(dout.reshape(B, T, 4, 16).transpose(1, 2))
        for i, t in enumerate((q, k, v)):
            assert _rel(got[:, i * 64:(i + 1) * 64].reshape(B, T, 4, 16).transpose(1, 2), t.grad) < 1.5e-2, i


def test_attention_dropout_is_consistent():
    """With dropout the three kernels must agree on the mask: finite differences of the forward output against the
    analytic gradients, and the kept fraction."""
    from attackfl_b200.ops.native import cuda_lib, ptr, check
    import ctypes
    B, T, p = 1, 100, 0.3
    lib = cuda_lib()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    qkv = torch.randn(B * T, 192, device="cuda")
    dout = torch.randn(B * T, 64, device="cuda")

    def fwd(t):
        o = torch.empty(B * T, 64, device="cuda"); l = torch.zeros(B * 4 * lib.afl_har_lse_stride(T), device="cuda")
        check(lib, lib.afl_har_attn_fwd(ptr(t), B, T, ctypes.c_float(p), ctypes.c_uint64(5), 1, 2, 3, ptr(o), ptr(l), st))
        return o, l
    out, lse = fwd(qkv)
    dqkv = torch.zeros_like(qkv)
    check(lib, lib.afl_har_attn_bwd(ptr(qkv), ptr(out), ptr(dout), ptr(lse), B, T, ctypes.c_float(p), ctypes.c_uint64(5), 1, 2,
                                    3, ptr(dqkv), st))
    g = torch.Generator(device="cuda").manual_seed(1)
    for _ in range(6):
        d = torch.randn(qkv.shape, device="cuda", generator=g)
        eps = 1e-2
        fd = ((fwd(qkv + eps * d)[0] - fwd(qkv - eps * d)[0]) * dout).sum() / (2 * eps)
        an = (dqkv * d).sum()
        assert abs(fd.item() - an.item()) < 2e-2 * max(1.0, abs(an.item())), (fd.item(), an.item())
    # v = 1 => output = kept mass / (1 - p): mean 1, so the kept fraction is ~ 1 - p
    ones = qkv.clone(); ones[:, 128:] = 1.0
    assert abs(fwd(ones)[0].mean().item() - 1.0) < 0.05


def test_forward_logits_match_module():
    mod, lay, row, tr, x, y = _setup()
    mod.eval()
    with torch.no_grad():
        want = mod(x[:, None, :])
    got = tr.logits_of(row, x)
    assert _rel(got, want) < 5e-3


def test_grad_step_matches_autograd():
    mod, lay, row, tr, x, y = _setup(B=5, seed=1)
    loss = torch.nn.functional.cross_entropy(mod(x[:, None, :]), y)
    loss.backward()
    want = {k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in mod.named_parameters()}
    grad = torch.zeros_like(row)
    tr.out3.zero_()
    out3 = tr.grad_step(row, grad, x, y, dropout=False)
    assert abs(out3[0].item() / 5 - loss.item()) < 2e-3 * max(1.0, abs(loss.item()))
    assert out3[1].item() == 0
    got = lay.views(grad)
    worst = {}
    for k, w in want.items():
        worst[k] = _rel(got[k], w)
    bad = {k: round(v, 4) for k, v in worst.items() if v > 2e-2}
    assert not bad, bad


def test_training_trajectory_matches_torch():
    """Six Adam steps (dropout off): the loss sequence of the native trainer follows PyTorch's."""
    from attackfl_b200.ops import flat
    mod, lay, row, tr, x, y = _setup(B=8, seed=2)
    opt = torch.optim.Adam(mod.parameters(), lr=2e-3)
    grad, m, v = torch.zeros_like(row), torch.zeros_like(row), torch.zeros_like(row)
    got, want = [], []
    for step in range(1, 7):
        tr.out3.zero_()
        tr.grad_step(row, grad, x, y, dropout=False)
        got.append(tr.out3[0].item() / 8)
        flat.adam_flat(row, grad, m, v, 2e-3, step)
        opt.zero_grad()
        loss = torch.nn.functional.cross_entropy(mod(x[:, None, :]), y)
        loss.backward()
        opt.step()
        want.append(loss.item())
    assert want[-1] < want[0]
    for a, b in zip(got, want):
        assert abs(a - b) < 3e-2 * max(1.0, abs(b)), (got, want)


def test_dropout_step_is_deterministic_per_seed():
    mod, lay, row, tr, x, y = _setup(B=4, seed=3)
    grad = torch.zeros_like(row)

    def run(seed, step):
        tr.out3.zero_()
        tr.grad_step(row, grad, x, y, seed=seed, client=1, step=step, dropout=True)
        return tr.out3[0].item(), grad.clone()
    l1, g1 = run(7, 1)
    l2, g2 = run(7, 1)
    l3, g3 = run(7, 2)
    assert abs(l1 - l2) < 1e-5 and torch.allclose(g1, g2, rtol=1e-3, atol=1e-6)     # atomics: summation order only
    assert abs(l1 - l3) > 1e-6 and not torch.allclose(g1, g3, rtol=1e-3, atol=1e-6)
    assert torch.isfinite(g1).all() and tr.out3[1].item() == 0


def test_engine_har_round_native_vs_torch(tmp_path):
    """One FL round of HAR through the engine: native trainer (default on CUDA) against backend: torch."""
    from attackfl_b200 import from_dict
    from attackfl_b200.engine import ClientSpec, Federation
    rows, metrics = [], []
    for backend in ("auto", "torch"):
        d = {"server": {"num-round": 1, "clients": 2, "mode": "fedavg", "model": "TransformerClassifier", "data-name": "HAR",
                        "data-distribution": {"num-data-range": [24, 32]}, "random-seed": 1},
             "learning": {"epoch": 1, "batch-size": 8, "learning-rate": 0.001, "dropout": False},
             "runtime": {"synthetic_train_rows": 48, "synthetic_test_rows": 32, "backend": backend}}
        fed = Federation(from_dict(d), [ClientSpec(0), ClientSpec(1)], device="cuda", workdir=str(tmp_path / backend),
                         verbose=False, save=False)
        assert (fed.har_trainer is not None) == (backend == "auto")
        res = fed.run()
        assert res[0].ok and 0.0 <= res[0].metric <= 1.0
        rows.append(fed.global_row.clone()); metrics.append(res[0].metric)
        fed.close()
    assert _rel(rows[0], rows[1]) < 2e-2


@pytest.mark.parametrize("mode", ["ffma", "tc-fwd"])
def test_grad_step_attention_fallback_modes(mode, monkeypatch):
    """ATTACKFL_HAR_ATTN selects the FFMA attention kernels (the path sequences longer than 568 tokens take) or the
    tensor-core forward with the FFMA backward; both must give the same gradients as autograd."""
    monkeypatch.setenv("ATTACKFL_HAR_ATTN", mode)
    mod, lay, row, tr, x, y = _setup(B=3, seed=5)
    assert tr.attn_tc == (mode != "ffma") and not (tr.attn_tc and tr.attn_bwd_tc)
    loss = torch.nn.functional.cross_entropy(mod(x[:, None, :]), y)
    loss.backward()
    grad = torch.zeros_like(row)
    tr.out3.zero_()
    tr.grad_step(row, grad, x, y, dropout=False)
    got = lay.views(grad)
    bad = {k: round(_rel(got[k], v.grad), 4) for k, v in mod.named_parameters() if _rel(got[k], v.grad) > 2e-2}
    assert not bad, bad


def test_partial_batch_and_dropout_masks_reach_backward():
    """A ragged last batch (B = 5 of max 8) with dropout: finite-difference check of the full step's loss against the
    analytic gradient along a random direction (same seed => same masks on every evaluation)."""
    mod, lay, row, tr, x, y = _setup(B=5, seed=6)
    grad = torch.zeros_like(row)

    def loss_at(r):
        tr.out3.zero_()
        tr.grad_step(r, grad, x, y, seed=11, client=2, step=4, dropout=True)
        return tr.out3[0].item() / 5
    base = loss_at(row)
    g0 = grad.clone()
    gen = torch.Generator(device="cuda").manual_seed(0)
    d = torch.randn(row.shape, device="cuda", generator=gen) * lay.valid_mask("cuda").float()
    d[lay.views(torch.arange(row.numel(), device="cuda").float())["pe.pe"].long().flatten()] = 0   # buffer, not trained
    eps = 2e-3
    fd = (loss_at(row + eps * d) - loss_at(row - eps * d)) / (2 * eps)
    an = float((g0 * d).sum())
    assert abs(fd - an) < 5e-2 * max(1.0, abs(an)), (fd, an, base)


def test_grad_step_batch_above_128():
    """Batches larger than the reference's 128 (the engine allows up to 512 on the native path).  With many samples a few
    head pre-activations land within TF32 noise of 0 and their ReLU derivative flips against the fp32 oracle (measured:
    4 of 8,192 at B = 128, which alone moves dpooled by 9 % in max-norm), so the check is on the loss and on the
    direction / size of the whole gradient; the per-tensor 2 % bound is tested at small batches above."""
    mod, lay, row, tr, x, y = _setup(B=130, seed=9)
    loss = torch.nn.functional.cross_entropy(mod(x[:, None, :]), y)
    loss.backward()
    grad = torch.zeros_like(row)
    tr.out3.zero_()
    tr.grad_step(row, grad, x, y, dropout=False)
    assert abs(tr.out3[0].item() / 130 - loss.item()) < 2e-3 * max(1.0, abs(loss.item()))
    want = lay.flatten({k: (v.grad if v.grad is not None else torch.zeros_like(v)) for k, v in
                        list(mod.named_parameters()) + [("pe.pe", mod.pe.pe)]}, device="cuda")
    cos = torch.nn.functional.cosine_similarity(grad, want, dim=0).item()
    assert cos > 0.995 and abs(grad.norm().item() / want.norm().item() - 1) < 0.05, cos
