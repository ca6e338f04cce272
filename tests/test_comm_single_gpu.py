"""``afl_fedavg_fused`` (csrc/comm.cu) on ONE GPU: the ranks of a 1- or 2-rank job are emulated by concurrent launches
on separate streams that are handed each other's buffers as "peer" pointers.  The flag protocol, the chunk pipeline, the
slice ownership and the LIE moments are exactly what runs across GPUs (only the NVLink hop and NVLS are missing), so the
kernel is covered by the single-GPU test run as well.  Reference: plain torch on the same rows."""
import ctypes

import pytest
import torch

from attackfl_b200.ops import native
from attackfl_b200.ops.native import check, ptr

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("world,R,A,D,chunks", [(1, 3, 0, 65536, 1), (2, 2, 2, 65536, 1), (1, 2, 1, 16 * 1024 * 1024, 0),
                                                (2, 1, 1, 262144, 4), (2, 3, 0, 1048576, 8)])
def test_fused_fedavg_with_emulated_ranks(world, R, A, D, chunks, monkeypatch):
    # both emulated ranks must be co-resident (cooperative grids that wait for each other): cap the blocks per launch;
    # `chunks` forces the pipeline depth so that the multi-chunk schedule is covered at small sizes too (0 = automatic)
    monkeypatch.setenv("ATTACKFL_FUSED_MAX_BLOCKS", "120")
    if chunks:
        monkeypatch.setenv("ATTACKFL_FUSED_CHUNKS", str(chunks))
    lib = native.cuda_lib()
    dev = "cuda"
    torch.manual_seed(world * 10 + A)
    C = world * R
    ref = torch.randn(D, device=dev) * 0.1
    rows = [(ref[None] + 0.05 * torch.randn(R, D, device=dev)).contiguous() for _ in range(world)]
    keep = torch.cat(rows).clone()
    wall = torch.rand(C) + 0.5
    wall = (wall / wall.sum()).to(dev)
    member_all = (torch.rand(max(A, 1), C) < 0.7).float()
    member_all[:, :2] = 1.0
    member_all = member_all.to(dev)
    kc = member_all.sum(1).contiguous()
    zz = torch.full((max(A, 1),), 0.74, device=dev)
    part = [torch.zeros(1 + 2 * A, D, device=dev) for _ in range(world)]
    glob = [torch.zeros(D, device=dev) for _ in range(world)]
    lie = [torch.zeros(max(A, 1), D, device=dev) for _ in range(world)]
    sig = [torch.zeros(2048, dtype=torch.int32, device=dev) for _ in range(world)]
    ctrl = [torch.zeros(128, dtype=torch.int32, device=dev) for _ in range(world)]
    PA = ctypes.c_void_p * world
    pp, gp, lp, sp = (PA(*[t.data_ptr() for t in ts]) for ts in (part, glob, lie, sig))
    streams = [torch.cuda.Stream() for _ in range(world)]
    torch.cuda.synchronize()
    for epoch in (1, 2):                       # two calls: the flags are monotonic epochs, nothing is reset in between
        for r in range(world):
            wl = wall[r * R:(r + 1) * R].contiguous()
            mem = member_all[:, r * R:(r + 1) * R].contiguous()
            with torch.cuda.stream(streams[r]):
                rc = lib.afl_fedavg_fused(ptr(rows[r]), ptr(wl), ctypes.c_int(R), ctypes.c_int64(D), ctypes.c_int(world),
                                          ctypes.c_int(r), pp, gp, lp if A else None, sp, ctypes.c_uint32(epoch), ctypes.c_int(A),
                                          ptr(ref) if A else None, ptr(mem) if A else None, ptr(kc) if A else None,
                                          ptr(zz) if A else None, ctypes.c_int(0), ptr(rows[r]) if epoch == 2 else None,
                                          ptr(ctrl[r]), ctypes.c_void_p(0), ctypes.c_void_p(0), ctypes.c_void_p(0),
                                          ctypes.c_void_p(streams[r].cuda_stream))
                check(lib, rc, "afl_fedavg_fused")
        torch.cuda.synchronize()
        want = (wall[:, None] * keep).sum(0)
        for r in range(world):
            assert int(ctrl[r][64].item()) == 0
            assert torch.allclose(glob[r], want, atol=2e-6, rtol=1e-5), (epoch, r, float((glob[r] - want).abs().max()))
            for q in range(A):
                sel = keep[member_all[q] > 0]
                want_l = sel.mean(0) + 0.74 * sel.std(0, unbiased=True)
                assert torch.allclose(lie[r][q], want_l, atol=2e-4), (epoch, r, q)
    for r in range(world):                     # the second call fanned the global model out to every local row
        assert torch.allclose(rows[r], want[None].expand(R, D), atol=2e-6, rtol=1e-5)
