#!/usr/bin/env python
"""BASELINE config 5: aggregation bandwidth sweep 1 MB - 1 GB at 2/4/8 GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench/agg_sweep.py

Every rank holds `--rows` client updates of S bytes (flat fp32 buffers of the stated size; a ~100 M-parameter
TransformerModel variant is 400 MB).  Timed per size, CUDA events, max over ranks, 3 warm-ups, 10 iterations:
  fused : K1-fused (csrc/comm.cu): weighted reduce of ALL ranks' rows read in place over NVLink + result written to
          every rank + fan-out to the local rows, one kernel per rank
  nccl  : the NCCL-only baseline: local weighted sum (K1 single-GPU kernel) + dist.all_reduce + local fan-out copy
Reported: time, algorithmic GB/s (S / t), all-reduce bus GB/s (2 (G-1)/G S / t) and the fraction of the roofline
max(NVLink bytes / 770 GB/s per direction, local HBM bytes / measured HBM) from B200_PROFILING.md.
"""
import argparse
import ctypes
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from attackfl_b200.ops import flat, native  # noqa: E402
from attackfl_b200.ops.native import check, ptr, stream_ptr  # noqa: E402
from attackfl_b200.parallel import ctx as pctx  # noqa: E402

NVLINK_GBS = 770.0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1, help="client rows per rank")
    ap.add_argument("--sizes-mb", default="1,4,16,64,256,1024")
    ap.add_argument("--iters", type=int, default=10)
    ap.add_argument("--tune", action="store_true",
                    help="time the fused kernel over pipeline depths (ATTACKFL_FUSED_CHUNKS) and grid caps (ATTACKFL_FUSED_MAX_BLOCKS)")
    a = ap.parse_args()
    ctx = pctx.init_from_env("cuda")
    assert ctx.distributed, "run under torchrun with >= 2 ranks"
    import torch.distributed._symmetric_memory as symm_mem
    G, rank, dev = ctx.world, ctx.rank, ctx.device
    hbm = 6540.0
    pk = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")
    if os.path.exists(pk):
        hbm = json.load(open(pk))["hbm_gbs"]
    lib = native.cuda_lib()
    R = a.rows
    C = R * G
    results = []
    for mb in [int(x) for x in a.sizes_mb.split(",")]:
        D = mb * 1024 * 1024 // 4
        D = D // (4 * G * 64) * (4 * G * 64)
        rows = symm_mem.empty(R, D, dtype=torch.float32, device=dev)
        glob = symm_mem.empty(D, dtype=torch.float32, device=dev)
        sig = symm_mem.empty(2048, dtype=torch.int32, device=dev)
        partial = symm_mem.empty(D, dtype=torch.float32, device=dev)
        rows.normal_(); sig.zero_()
        h_rows, h_glob, h_sig, h_part = (symm_mem.rendezvous(t, dist.group.WORLD) for t in (rows, glob, sig, partial))
        torch.cuda.synchronize(); dist.barrier()
        PA = ctypes.c_void_p * G
        gp, sp = PA(*[int(p) for p in h_glob.buffer_ptrs]), PA(*[int(p) for p in h_sig.buffer_ptrs])
        pp = PA(*[int(p) for p in h_part.buffer_ptrs])
        ctrl = torch.zeros(128, dtype=torch.int32, device=dev)
        w = torch.full((C,), 1.0 / C, device=dev)
        wl = torch.full((R,), 1.0 / C, device=dev)
        epoch = [0]

        mc_part, mc_glob = int(getattr(h_part, "multicast_ptr", 0) or 0), int(getattr(h_glob, "multicast_ptr", 0) or 0)

        def fused(nvls=False):
            epoch[0] += 1
            check(lib, lib.afl_fedavg_fused(ptr(rows), ptr(wl), ctypes.c_int(R), ctypes.c_int64(D), ctypes.c_int(G),
                                            ctypes.c_int(rank), pp, gp, None, sp, ctypes.c_uint32(epoch[0]), ctypes.c_int(0),
                                            None, None, None, None, ctypes.c_int(0), ptr(rows), ptr(ctrl),
                                            ctypes.c_void_p(mc_part if nvls else 0), ctypes.c_void_p(mc_glob if nvls else 0),
                                            ctypes.c_void_p(0), stream_ptr(rows)), "afl_fedavg_fused")

        part = torch.empty(D, device=dev)

        def nccl():
            flat.fedavg(rows, wl, out=part)
            dist.all_reduce(part)
            rows.copy_(part[None].expand(R, D))

        def timeit(fn):
            for _ in range(3):
                fn()
            torch.cuda.synchronize(); dist.barrier()
            ts = []
            for _ in range(a.iters):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                dist.barrier()
                e0.record(); fn(); e1.record()
                torch.cuda.synchronize()
                ts.append(e0.elapsed_time(e1))
            t = torch.tensor([sorted(ts)[len(ts) // 2]], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return float(t.item()) * 1e-3

        # correctness: both paths give the same global model
        rows.normal_(); torch.cuda.synchronize(); dist.barrier()
        keep = rows.clone()
        fused(); torch.cuda.synchronize(); dist.barrier()
        g1 = glob.clone(); rows.copy_(keep); torch.cuda.synchronize(); dist.barrier()
        nccl(); torch.cuda.synchronize()
        err = float((g1 - part).abs().max())
        rows.copy_(keep); torch.cuda.synchronize(); dist.barrier()
        S = D * 4
        paths = [("fused", fused), ("nccl", nccl)]
        if mc_part and mc_glob:
            rows.copy_(keep); torch.cuda.synchronize(); dist.barrier()
            fused(True); torch.cuda.synchronize(); dist.barrier()
            err_nvls = float((glob - part).abs().max())
            rows.copy_(keep); torch.cuda.synchronize(); dist.barrier()
            paths.insert(1, ("fused_nvls", lambda: fused(True)))
        if a.tune:
            best = None
            fn = (lambda: fused(True)) if (mc_part and mc_glob) else fused
            for blocks in (148, 296, 592):
                for chunks in (1, 2, 4, 8, 16, 32):
                    os.environ["ATTACKFL_FUSED_CHUNKS"] = str(chunks)
                    os.environ["ATTACKFL_FUSED_MAX_BLOCKS"] = str(blocks)
                    t = timeit(fn)
                    rec = {"tune": True, "size_mb": mb, "gpus": G, "rows_per_rank": R, "chunks": chunks, "max_blocks": blocks,
                           "nvls": bool(mc_part and mc_glob), "ms": round(t * 1e3, 4)}
                    if best is None or t < best[0]:
                        best = (t, chunks, blocks)
                    if rank == 0:
                        print(json.dumps(rec), flush=True)
            os.environ.pop("ATTACKFL_FUSED_CHUNKS", None); os.environ.pop("ATTACKFL_FUSED_MAX_BLOCKS", None)
            t = timeit(fn)
            if rank == 0:
                print(json.dumps({"tune": True, "size_mb": mb, "default_ms": round(t * 1e3, 4), "best_ms": round(best[0] * 1e3, 4),
                                  "best_chunks": best[1], "best_max_blocks": best[2]}), flush=True)
            paths = []
        for name, fn in paths:
            t = timeit(fn)
            nv_bytes = (G - 1) / G * S              # per direction per rank: partial slices in, result slices out
            hbm_bytes = (R * S + S) + (S / G + S) + (S + R * S)     # pre-reduce, slice reduce + result, fan-out
            roof = max(nv_bytes / (NVLINK_GBS * 1e9), hbm_bytes / (hbm * 1e9))
            rec = {"size_mb": mb, "gpus": G, "rows_per_rank": R, "path": name, "ms": round(t * 1e3, 4),
                   "alg_GBps": round(S / t / 1e9, 1), "bus_GBps": round(2 * (G - 1) / G * S / t / 1e9, 1),
                   "roofline_frac": round(roof / t, 3),
                   "max_abs_diff_vs_nccl_path": err_nvls if name == "fused_nvls" else err}
            results.append(rec)
            if rank == 0:
                print(json.dumps(rec), flush=True)
        if int(ctrl[64].item()) != 0:
            raise RuntimeError("fused collective timed out")
        del rows, glob, sig, part, keep, g1, partial
        torch.cuda.empty_cache()
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
