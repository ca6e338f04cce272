#!/usr/bin/env python
"""Per-kernel micro-benchmarks with roofline fractions (1 GPU): K1 FedAvg(+moments), K3 Gram (tcgen05), K4 craft,
K5a/K5b hyper emit/step, K6 select.  CUDA events, 3 warm-ups, L2 flushed between iterations (256 MiB write).
Peaks come from MEASURED_PEAKS.json (fallback 6.65 TB/s / 1.59 PFLOP/s, B200_PROFILING.md).

    python bench/kernels.py [--only gram,fedavg,...] [--json out.json]
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from attackfl_b200 import FlatLayout, models  # noqa: E402
from attackfl_b200.models.hyper import FusedHyperNet  # noqa: E402
from attackfl_b200.ops import flat, gram_native  # noqa: E402


def peaks():
    p = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return d["hbm_gbs"], d["bf16_tflops"], "measured"
    return 6650.0, 1590.0, "fallback"


FLUSH = None


def timeit(fn, iters=10, warm=3):
    global FLUSH
    if FLUSH is None:
        FLUSH = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device="cuda")
    for _ in range(warm):
        fn()
    ts = []
    for _ in range(iters):
        FLUSH.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2] * 1e-3          # median seconds


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--only", default="")
    ap.add_argument("--json", default="")
    args = ap.parse_args()
    only = set(x for x in args.only.split(",") if x)
    hbm, tf, src = peaks()
    out = {"peaks": {"hbm_gbs": hbm, "bf16_tflops": tf, "source": src}, "rows": []}
    dev = "cuda"

    def rec(kernel, shape, sec, bytes_=None, flops=None, note=""):
        r = {"kernel": kernel, "shape": shape, "us": round(sec * 1e6, 2), "note": note}
        if bytes_:
            r["GBps"] = round(bytes_ / sec / 1e9, 1); r["hbm_frac_of_" + src] = round(bytes_ / sec / 1e9 / hbm, 3)
        if flops:
            r["TFLOPs"] = round(flops / sec / 1e12, 2); r["tensor_frac_of_" + src] = round(flops / sec / 1e12 / tf, 3)
        out["rows"].append(r)
        print(json.dumps(r), flush=True)

    if not only or "fedavg" in only:
        for C, D in ((8, 48128), (64, 48128), (128, 98304), (8, 25_000_000), (16, 25_000_000), (4, 100_000_000 // 4 * 4)):
            rows = torch.randn(C, D, device=dev)
            w = torch.full((C,), 1.0 / C, device=dev)
            o = torch.empty(D, device=dev)
            s = timeit(lambda: flat.fedavg(rows, w, out=o))
            rec("K1 fedavg (reduce+scale)", f"C={C} D={D}", s, bytes_=(C + 1) * D * 4)
            s = timeit(lambda: flat.fedavg(rows, w, out=o, bcast=rows))
            rec("K1 fedavg (+broadcast to C rows)", f"C={C} D={D}", s, bytes_=(2 * C + 1) * D * 4)
            if D < 1_000_000:
                mem = (torch.rand(2, C, device=dev) < 0.6).float(); mem[:, :2] = 1
                z = torch.tensor([0.74, 0.74], device=dev)
                s = timeit(lambda: flat.fedavg(rows, w, out=o, member=mem, z=z, ref=o, want_lie=True))
                rec("K1+K2 fedavg + 2 LIE moment sets", f"C={C} D={D}", s, bytes_=(C + 1 + 6) * D * 4)
            del rows
    if not only or "gram" in only:
        for R, D in ((18, 48128), (66, 98304), (130, 98304), (256, 203776), (66, 25_000_000), (130, 25_000_000)):
            a = torch.randn(R, D, device=dev) * 0.05
            rb = 128 if R <= 128 else 256
            ncol = (R + 15) // 16 * 16
            gram_native.gram_tcgen05(a, "tf32")
            s_tf = timeit(lambda: gram_native.gram_tcgen05(a, "tf32"))
            rec("K3 gram tcgen05 tf32 (TMA on fp32 rows + mma + reduce)", f"R={R} D={D}", s_tf, bytes_=R * D * 4,
                flops=2.0 * rb * ncol * D, note="one pass over the fp32 rows; FLOPs on the padded tile actually issued")
            gram_native.gram_tcgen05(a, "bf16")
            s_all = timeit(lambda: gram_native.gram_tcgen05(a, "bf16"))
            rec("K3 gram tcgen05 bf16 (cast+mma+reduce)", f"R={R} D={D}", s_all, bytes_=R * D * 4 + rb * D * 2 * 2,
                flops=2.0 * rb * ncol * D, note="FLOPs counted on the padded 128/256 x N tile actually issued")
            s_t = timeit(lambda: (a @ a.t()))
            rec("torch fp32 matmul A A^T (library baseline)", f"R={R} D={D}", s_t, bytes_=R * D * 4, flops=2.0 * R * R * D)
            del a
    if not only or "select" in only:
        for C in (8, 64, 128):
            rows = torch.randn(C, 98304, device=dev)
            s = timeit(lambda: flat.coord_select(rows, "median"))
            rec("K6 median", f"C={C} D=98304", s, bytes_=(C + 1) * 98304 * 4)
            s = timeit(lambda: torch.median(rows, dim=0).values)
            rec("torch.median (library baseline)", f"C={C} D=98304", s, bytes_=(C + 1) * 98304 * 4)
    if not only or "hyper" in only:
        for name, n in (("TransformerModel", 8), ("CNNModel", 8), ("RNNModel", 64)):
            torch.manual_seed(0)
            h = models.HyperNetwork(models.build(name), n, 8, 100, False, 2)
            fused = FusedHyperNet.from_module(h, device=dev)
            nat = flat.NativeHyper(fused)
            D = fused.layout.padded
            theta = nat.emit(list(range(n)))
            s = timeit(lambda: nat.emit(list(range(n)), out=theta))
            rec("K5a hyper emit", f"{name} n={n} D={D}", s, bytes_=D * 100 * 4 + n * D * 4, flops=2.0 * n * D * 100,
                note="FFMA kernel; tensor fraction shown for reference only")
            row = theta[0].clone()
            s = timeit(lambda: nat.step(0, row, 1e-3, 1.0))
            rec("K5b hyper step (1 client)", f"{name} D={D}", s, bytes_=D * 100 * 4 * 7 + D * 4 * 8,
                note="W_h read once for delta, then W,m,v read+written by Adam")
    if args.json:
        json.dump(out, open(args.json, "w"), indent=1)


if __name__ == "__main__":
    main()
