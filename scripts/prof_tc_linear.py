import sys, torch
sys.path.insert(0, ".")
from attackfl_b200.ops import tc_linear as tc
M = 128 * 561
x = torch.randn(M, 64, device="cuda"); w = torch.randn(192, 64, device="cuda") / 8; b = torch.randn(192, device="cuda")
y = torch.empty(M, 192, device="cuda"); dy = torch.randn(M, 192, device="cuda"); dw = torch.empty(192, 64, device="cuda"); db = torch.empty(192, device="cuda")
for _ in range(3):
    tc.linear_fwd(x, w, b, out=y)
    tc.linear_dw(dy, x, dw=dw, db=db)
torch.cuda.synchronize()
print("ok")
