import torch, sys
sys.path.insert(0, ".")
from attackfl_b200.ops import tc_linear as tc
torch.manual_seed(0)
def rel(a, b): return ((a - b).abs().max() / (b.abs().max() + 1e-12)).item()
for (m, n, k) in [(128, 64, 64), (128, 32, 32), (256, 192, 64), (1000, 64, 256)]:
    x = torch.randn(m, k, device="cuda"); w = torch.randn(n, k, device="cuda") / k ** 0.5; dy = torch.randn(m, n, device="cuda")
    got = tc.linear_dx(dy, w); want = dy @ w
    print(f"dx m={m} n={n} k={k}: rel={rel(got, want):.3e} nz={int((got != 0).sum())}/{got.numel()}")
    if rel(got, want) > 1e-2:
        print("   got ", got[0, :8].tolist()); print("   want", want[0, :8].tolist())
        # hypotheses: columns permuted? transposed weight?
        if n == k:
            print("   vs dy@w.T", rel(got, dy @ w.t()))
    dw, db = tc.linear_dw(dy, x); wantw = dy.t() @ x
    print(f"dw m={m} n={n} k={k}: rel={rel(dw, wantw):.3e} nz={int((dw != 0).sum())}/{dw.numel()} db rel={rel(db, dy.sum(0)):.3e}")
    if rel(dw, wantw) > 1e-2:
        print("   got ", dw[0, :8].tolist()); print("   want", wantw[0, :8].tolist())
        print("   vs transposed", rel(dw, wantw.t()) if n == k else None)
