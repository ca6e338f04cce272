#!/usr/bin/env python
"""HAR linear-layer shapes (M = 128 x 561 tokens): tcgen05 kernels vs torch (cuBLAS fp32 and TF32) — CUDA events,
256 MiB L2 flush between timed calls."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from attackfl_b200.ops import tc_linear as tc  # noqa: E402


def timeit(fn, iters=20):
    flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
    for _ in range(3):
        fn()
    ts = []
    for _ in range(iters):
        flush.zero_()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b) * 1e3)
    ts.sort()
    return ts[len(ts) // 2]


def main():
    M = 128 * 561
    for name, n, k, relu in (("in_proj", 192, 64, False), ("out_proj", 64, 64, False), ("ffn1+relu", 256, 64, True),
                             ("ffn2", 64, 256, False)):
        x = torch.randn(M, k, device="cuda"); w = torch.randn(n, k, device="cuda") / k ** 0.5
        b = torch.randn(n, device="cuda"); dy = torch.randn(M, n, device="cuda")
        y = torch.empty(M, n, device="cuda"); dx = torch.empty(M, k, device="cuda")
        dw = torch.empty(n, k, device="cuda"); db = torch.empty(n, device="cuda")
        rec = {"layer": name, "M": M, "N": n, "K": k}
        rec["fwd_us"] = timeit(lambda: tc.linear_fwd(x, w, b, relu=relu, out=y))
        rec["dx_us"] = timeit(lambda: tc.linear_dx(dy, w, out=dx))
        rec["dw_us"] = timeit(lambda: tc.linear_dw(dy, x, dw=dw, db=db))
        for tf32 in (False, True):
            torch.backends.cuda.matmul.allow_tf32 = tf32
            tag = "tf32" if tf32 else "fp32"
            f = (lambda: torch.relu(torch.addmm(b, x, w.t()))) if relu else (lambda: torch.addmm(b, x, w.t()))
            rec[f"torch_{tag}_fwd_us"] = timeit(f)
            rec[f"torch_{tag}_dx_us"] = timeit(lambda: dy @ w)
            rec[f"torch_{tag}_dw_us"] = timeit(lambda: (dy.t() @ x, dy.sum(0)))
        torch.backends.cuda.matmul.allow_tf32 = False
        bytes_f = 4 * (M * k + M * n + n * k)
        rec["fwd_GBps"] = round(bytes_f / rec["fwd_us"] / 1e3, 1)
        rec["dw_GBps"] = round(4 * (M * k + M * n) / rec["dw_us"] / 1e3, 1)
        print(json.dumps({k_: (round(v, 2) if isinstance(v, float) else v) for k_, v in rec.items()}), flush=True)


if __name__ == "__main__":
    main()
