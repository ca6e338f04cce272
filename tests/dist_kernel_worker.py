"""Kernel-level multi-rank worker: ``afl_fedavg_fused`` (csrc/comm.cu) against plain torch math on NCCL-gathered rows.

    python -m torch.distributed.run --nproc-per-node N tests/dist_kernel_worker.py OUT [stress_epochs]

Covers one-chunk and multi-chunk pipelines, 1 / 3 rows per rank, LIE moment sets, P2P and NVLS (when the symmetric
buffers have a multicast mapping), and a flag-protocol stress with random per-rank device delays.  Full vectors are compared.
"""
import ctypes
import json
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from attackfl_b200.ops import native  # noqa: E402
from attackfl_b200.ops.native import check, ptr, stream_ptr  # noqa: E402
from attackfl_b200.parallel import ctx as pctx  # noqa: E402


def main():
    out = sys.argv[1]
    stress = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    ctx = pctx.init_from_env("cuda")
    import torch.distributed._symmetric_memory as symm_mem
    G, rank, dev = ctx.world, ctx.rank, ctx.device
    lib = native.cuda_lib()
    rec = {"world": G, "cases": [], "stress": None}
    sig = symm_mem.empty(2048, dtype=torch.int32, device=dev)
    sig.zero_()
    h_sig = symm_mem.rendezvous(sig, dist.group.WORLD)
    PA = ctypes.c_void_p * G
    sp = PA(*[int(p) for p in h_sig.buffer_ptrs])
    ctrl = torch.zeros(128, dtype=torch.int32, device=dev)
    epoch = [0]

    def case(D, R, A, nvls, iters=1, jitter=None):
        rows = symm_mem.empty(R, D, dtype=torch.float32, device=dev)
        glob = symm_mem.empty(D, dtype=torch.float32, device=dev)
        part = symm_mem.empty(1 + 2 * A, D, dtype=torch.float32, device=dev)
        lie = symm_mem.empty(max(A, 1), D, dtype=torch.float32, device=dev)
        hs = [symm_mem.rendezvous(t, dist.group.WORLD) for t in (rows, glob, part, lie)]
        torch.cuda.synchronize(); dist.barrier()
        gp, pp, lp = (PA(*[int(p) for p in h.buffer_ptrs]) for h in (hs[1], hs[2], hs[3]))
        mc = [int(getattr(h, "multicast_ptr", 0) or 0) for h in (hs[2], hs[1], hs[3])]
        if nvls and not all(mc):
            return None
        g = torch.Generator(device=dev).manual_seed(1000 + rank)
        C = R * G
        wall = torch.rand(C, generator=torch.Generator().manual_seed(5)) + 0.5
        wall = (wall / wall.sum()).to(dev)
        wl = wall[rank * R:(rank + 1) * R].contiguous()
        ref = torch.randn(D, generator=torch.Generator().manual_seed(6)).to(dev) * 0.1
        member_all = (torch.rand(max(A, 1), C, generator=torch.Generator().manual_seed(7)) < 0.7).float()
        member_all[:, :2] = 1.0
        member_all = member_all.to(dev)
        kc = member_all.sum(1).contiguous()
        mem = member_all[:, rank * R:(rank + 1) * R].contiguous()
        zz = torch.full((max(A, 1),), 0.74, device=dev)
        worst = 0.0
        for it in range(iters):
            rows.normal_(generator=g)
            rows.mul_(0.05).add_(ref[None])
            torch.cuda.synchronize(); dist.barrier()
            allrows = torch.empty(C, D, device=dev)
            dist.all_gather_into_tensor(allrows, rows.clone())
            want = (wall[:, None] * allrows).sum(0)
            if jitter is not None:
                torch.cuda._sleep(jitter.randrange(0, 3_000_000))
            epoch[0] += 1
            rc = lib.afl_fedavg_fused(ptr(rows), ptr(wl), ctypes.c_int(R), ctypes.c_int64(D), ctypes.c_int(G), ctypes.c_int(rank),
                                      pp, gp, lp if A else None, sp, ctypes.c_uint32(epoch[0]), ctypes.c_int(A),
                                      ptr(ref) if A else None, ptr(mem) if A else None, ptr(kc) if A else None,
                                      ptr(zz) if A else None, ctypes.c_int(0), ptr(rows), ptr(ctrl),
                                      ctypes.c_void_p(mc[0] if nvls else 0), ctypes.c_void_p(mc[1] if nvls else 0),
                                      ctypes.c_void_p(mc[2] if (nvls and A) else 0), stream_ptr(rows))
            check(lib, rc, "afl_fedavg_fused")
            torch.cuda.synchronize()
            err = float((glob - want).abs().max())
            err = max(err, float((rows - want[None]).abs().max()))          # fan-out reached every local row
            if A:
                for q in range(A):
                    sel = allrows[member_all[q] > 0]
                    want_l = sel.mean(0) + 0.74 * sel.std(0, unbiased=True)
                    err = max(err, float((lie[q] - want_l).abs().max()))
            worst = max(worst, err)
            dist.barrier()
        assert int(ctrl[64].item()) == 0, "fused kernel reported a time-out"
        del rows, glob, part, lie
        return worst

    for D in (1024 * G, 1 << 20, 1 << 24):
        for R in (1, 3):
            for A in (0, 2):
                if D == (1 << 24) and (R > 1 or A > 0):
                    continue
                for nvls in (False, True):
                    e = case(D, R, A, nvls)
                    if e is not None:
                        rec["cases"].append({"D": D, "R": R, "A": A, "nvls": nvls, "max_abs_err": e})
    if stress:
        rng = random.Random(99 + 31 * rank)
        e = case(2048 * G, 2, 1, False, iters=stress, jitter=rng)
        e2 = case(1 << 22, 1, 0, G >= 4, iters=max(stress // 50, 2), jitter=rng)       # multi-chunk pipeline under skew
        rec["stress"] = {"epochs": stress, "max_abs_err": e, "multi_chunk_max_abs_err": e2}
    if rank == 0:
        with open(out, "w") as fh:
            json.dump(rec, fh)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
