"""Tensor-core trainer timing: C clients x (n / 128) optimiser steps in one launch, CUDA events.
Usage: python scripts/prof_train_tc.py [n_data] [epochs]   (PROF_CLIENTS=8)"""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from attackfl_b200 import FlatLayout, models
from attackfl_b200.data import make_round_indices, synthetic_icu
from attackfl_b200.ops import train_tc as TC

name = "TransformerModel"
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1280
epochs = int(sys.argv[2]) if len(sys.argv) > 2 else 1
dev = "cuda"
for C in [int(x) for x in os.environ.get("PROF_CLIENTS", "8").split(",")]:
    torch.manual_seed(0)
    lay = FlatLayout.from_module(models.build(name))
    rows = torch.stack([lay.flatten(models.build(name).state_dict()) for _ in range(C)]).to(dev)
    m, v = torch.zeros_like(rows), torch.zeros_like(rows)
    gw = torch.zeros_like(rows)
    stash = torch.zeros(C, TC.stash_floats_per_client(dev), device=dev)
    data = synthetic_icu(16384, 1).packed().to(dev)
    g = torch.Generator(device=dev).manual_seed(0)
    idx = make_round_indices(16384, [n] * C, epochs, g, dev)
    nd = torch.full((C,), n, dtype=torch.int32, device=dev)
    cid = torch.arange(C, dtype=torch.int32, device=dev)
    for it in range(3):
        m.zero_(); v.zero_()
        torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
        e0.record()
        loss, steps = TC.train_round_tc(lay, rows, m, v, gw, stash, data, idx, nd, cid, 128, epochs, seed=it, dropout=True)
        e1.record(); torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        st = int(steps[0])
        print(f"C={C} cluster={TC.last_cluster(dev)} tc train launch {ms:.3f} ms, {st} steps/client, {1e3 * ms / max(st,1):.1f} us/step, loss {float(loss[0,0])/n/epochs:.4f}")
if os.environ.get("PROF_STAMPS"):
    import ctypes
    from attackfl_b200.ops import native
    lib = native.cuda_lib()
    dbg = torch.zeros(64, dtype=torch.int64, device=dev)
    lib.afl_tc_set_debug(ctypes.c_void_p(dbg.data_ptr()))
    m.zero_(); v.zero_()
    TC.train_round_tc(lay, rows, m, v, gw, stash, data, idx, nd, cid, 128, epochs, seed=1, dropout=True)
    torch.cuda.synchronize()
    lib.afl_tc_set_debug(ctypes.c_void_p(0))
    t = dbg.cpu().tolist()
    names = ["E0", "E1", "E2", "E3", "E4", "E5", "E6", "E7", "E8", "B7", "B6", "B5", "B4", "B3", "B2", "B1", "sg->gw", "adam", "load_small", "end"]
    k = max(i for i, x in enumerate(t) if x)
    print("stamps (cycles):", " ".join(f"{names[i] if i < len(names) else i}:{t[i + 1] - t[i]}" for i in range(k)), "total", t[k] - t[0])
    print("E3 detail:", [t[32 + i + 1] - t[32 + i] for i in range(6)], "abs0", t[32] - t[0])
    print("producer P3(V): start", t[48] - t[0], "wait_a", t[49] - t[48], "wait_w", t[50] - t[49], "issue", t[51] - t[50], "; E3 TS2(0) at", t[32] - t[0], "TS2(5) at", t[37] - t[0])
    print("optimiser hand-over (main): fence+arrives", t[41] - t[40], "wait helpers", t[43] - t[41])
    print("optimiser hand-over (helper 1): early for", t[53] - t[52], "grad load+update+stores", t[54] - t[53], "barrier", t[55] - t[54], "fence+arrive", t[56] - t[55])
