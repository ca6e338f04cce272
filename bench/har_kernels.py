#!/usr/bin/env python
"""Per-kernel timing of the native HAR step at batch 128 (CUDA events, median of 10)."""
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from attackfl_b200.ops import tc_linear as tc  # noqa: E402
from attackfl_b200.ops.native import check, cuda_lib, ptr  # noqa: E402

c_f, c_u64, c_i64 = ctypes.c_float, ctypes.c_uint64, ctypes.c_int64


def timeit(fn, iters=10):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return round(ts[len(ts) // 2], 1)


def main():
    B, T = int(os.environ.get("HAR_B", "128")), 561
    M = B * T
    lib = cuda_lib()
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    z = lambda *s: torch.randn(*s, device="cuda")                  # noqa: E731
    qkv, att, dout, lse, dqkv = z(M, 192), z(M, 64), z(M, 64), torch.zeros(B * 4 * lib.afl_har_lse_stride(T), device="cuda"), z(M, 192)
    mask = torch.zeros(B * 4 * lib.afl_har_attn_mask_words(T) * lib.afl_har_lse_stride(T), dtype=torch.int32, device="cuda")
    delta = torch.zeros(B * 4 * lib.afl_har_lse_stride(T), device="cuda")
    x64, r64, y64, xh, rs = z(M, 64), z(M, 64), z(M, 64), z(M, 64), torch.rand(M, device="cuda")
    ga, be, dg, db = z(64), z(64), torch.zeros(64, device="cuda"), torch.zeros(64, device="cuda")
    out = {"batch": B}
    for p in (0.0, 0.1):
        tag = f"p{p}"
        out[f"attn_fwd_{tag}_us"] = timeit(lambda: check(lib, lib.afl_har_attn_fwd(ptr(qkv), B, T, c_f(p), c_u64(1), 0, 1, 0, ptr(att), ptr(lse), st)))
        out[f"attn_fwd_tc_{tag}_us"] = timeit(lambda: check(lib, lib.afl_har_attn_fwd_tc(ptr(qkv), B, T, c_f(p), c_u64(1), 0, 1, 0, ptr(att), ptr(lse), ptr(mask), st)))
        out[f"attn_bwd_{tag}_us"] = timeit(lambda: check(lib, lib.afl_har_attn_bwd(ptr(qkv), ptr(att), ptr(dout), ptr(lse), B, T, c_f(p), c_u64(1), 0, 1, 0, ptr(dqkv), st)))
        out[f"attn_bwd_tc_{tag}_us"] = timeit(lambda: check(lib, lib.afl_har_attn_bwd_tc(ptr(qkv), ptr(att), ptr(dout), ptr(lse), ptr(mask), B, T, c_f(p), ptr(delta), ptr(dqkv), st)))
        out[f"add_ln_fwd_{tag}_us"] = timeit(lambda: check(lib, lib.afl_har_add_ln_fwd(ptr(x64), ptr(r64), c_f(p), c_u64(1), 0, 1, 8, ptr(ga), ptr(be), c_i64(M), ptr(y64), ptr(xh), ptr(rs), st)))
        out[f"ln_bwd_{tag}_us"] = timeit(lambda: check(lib, lib.afl_har_ln_bwd(ptr(x64), None, T, c_f(1.0), ptr(xh), ptr(rs), ptr(ga), c_f(p), c_u64(1), 0, 1, 8, c_i64(M), ptr(y64), ptr(r64), ptr(dg), ptr(db), st)))
    xin, cw, cb, pe = z(B, T), z(64, 1, 3), z(64), z(1, 600, 64)
    out["embed_fwd_us"] = timeit(lambda: check(lib, lib.afl_har_embed_fwd(ptr(xin), ptr(cw), ptr(cb), ptr(pe), B, T, ptr(x64), st)))
    dcw, dcb = torch.zeros(64, 1, 3, device="cuda"), torch.zeros(64, device="cuda")
    out["embed_bwd_us"] = timeit(lambda: check(lib, lib.afl_har_embed_bwd(ptr(x64), ptr(xin), B, T, ptr(dcw), ptr(dcb), st)))
    pooled = torch.empty(B, 64, device="cuda")
    out["pool_fwd_us"] = timeit(lambda: check(lib, lib.afl_har_pool_fwd(ptr(x64), B, T, ptr(pooled), st)))
    q, k, v = (qkv[:, i * 64:(i + 1) * 64].reshape(B, T, 4, 16).transpose(1, 2).contiguous().requires_grad_(True) for i in range(3))
    out["torch_sdpa_fwd_us"] = timeit(lambda: torch.nn.functional.scaled_dot_product_attention(q, k, v))

    def sdpa_fb():
        o = torch.nn.functional.scaled_dot_product_attention(q, k, v)
        o.backward(torch.ones_like(o))
    out["torch_sdpa_fwd_bwd_us"] = timeit(sdpa_fb)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
