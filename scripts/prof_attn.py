import ctypes, sys, torch
sys.path.insert(0, ".")
from attackfl_b200.ops.native import check, cuda_lib, ptr
lib = cuda_lib()
B, T = 128, 561
M = B * T
st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
tp, mw = lib.afl_har_lse_stride(T), lib.afl_har_attn_mask_words(T)
qkv = torch.randn(M, 192, device="cuda"); dout = torch.randn(M, 64, device="cuda")
out = torch.empty(M, 64, device="cuda"); lse = torch.zeros(B * 4 * tp, device="cuda")
mask = torch.zeros(B * 4 * mw * tp, dtype=torch.int32, device="cuda"); delta = torch.zeros(B * 4 * tp, device="cuda")
dqkv = torch.empty(M, 192, device="cuda")
p = ctypes.c_float(0.1)
for _ in range(2):
    check(lib, lib.afl_har_attn_fwd_tc(ptr(qkv), B, T, p, ctypes.c_uint64(1), 0, 1, 0, ptr(out), ptr(lse), ptr(mask), st))
    check(lib, lib.afl_har_attn_bwd_tc(ptr(qkv), ptr(out), ptr(dout), ptr(lse), ptr(mask), B, T, p, ptr(delta), ptr(dqkv), st))
torch.cuda.synchronize()
print("ok")
