import sys, torch
sys.path.insert(0, ".")
from attackfl_b200.ops import tc_linear as tc
torch.manual_seed(0)
def rel(a, b): return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()
def rowerr(a, b):
    e = (a - b).abs().amax(1) / (b.abs().max() + 1e-12)
    bad = (e > 5e-3).nonzero().flatten()
    return int(bad.numel()), bad[:8].tolist()
for (m, n, k) in [(71808, 192, 64), (71808, 64, 64), (71808, 256, 64), (71808, 64, 256), (40000, 64, 64)]:
    x = torch.randn(m, k, device="cuda"); w = torch.randn(n, k, device="cuda") / k ** 0.5; b = torch.randn(n, device="cuda")
    dy = torch.randn(m, n, device="cuda")
    y = tc.linear_fwd(x, w, b); want = x @ w.t() + b
    print(f"fwd m={m} n={n} k={k}: rel={rel(y, want):.2e} bad rows {rowerr(y, want)}")
    dx = tc.linear_dx(dy, w); want = dy @ w
    print(f"dx : rel={rel(dx, want):.2e} bad rows {rowerr(dx, want)}")
    acc = torch.randn(m, k, device="cuda"); want = acc + dy @ w
    got = tc.linear_dx(dy, w, out=acc.clone(), accumulate=True)
    print(f"dx+: rel={rel(got, want):.2e} bad rows {rowerr(got, want)}")
    dw, db = tc.linear_dw(dy, x); wantw = dy.t() @ x
    print(f"dw : rel={rel(dw, wantw):.2e} db rel={rel(db, dy.sum(0)):.2e}")
    if n == 64:
        resid, ga, be = torch.randn(m, 64, device="cuda"), torch.randn(64, device="cuda"), torch.randn(64, device="cuda")
        yy, xh, rs = torch.empty(m, 64, device="cuda"), torch.empty(m, 64, device="cuda"), torch.empty(m, device="cuda")
        tc.linear_ln_fwd(x, w, b, resid, ga, be, yy, xh, rs)
        want = torch.nn.functional.layer_norm(resid + x @ w.t() + b, (64,), ga, be)
        print(f"ln : rel={rel(yy, want):.2e} bad rows {rowerr(yy, want)}")
