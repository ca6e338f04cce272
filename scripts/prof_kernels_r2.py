"""One launch of every native kernel that had no ncu capture yet (round-1 verdict, item 8), at representative shapes, so
that a single `ncu --set full` run over this script yields a raw metrics row per kernel:
  K1 single-GPU (fedavg_kernel), K1-fused (fedavg_fused_kernel, world = 1: flags to itself), K6 (select_warp_kernel),
  one-Gram attack path (multi_moments / gram / multi_gamma / multi_craft), hyper round + emit (hyper_round_kernel,
  hyper_feat_kernel, hyper_emit_rows_kernel), ScionFL (scion_*), inference (infer_kernel), one HAR step (har_kernels.cu,
  gemm / attention tcgen05 kernels), the interpreter and the tensor-core trainer (train_kernel, tc_train_kernel).
Usage: python scripts/prof_kernels_r2.py"""
import ctypes
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch

from attackfl_b200 import FlatLayout, models
from attackfl_b200.data import make_round_indices, synthetic_icu
from attackfl_b200.models.hyper import FusedHyperNet, HyperNetwork
from attackfl_b200.models.programs import build_program
from attackfl_b200.ops import flat, native, train as T, train_tc as TC
from attackfl_b200.ops.native import check, ptr, stream_ptr

dev = "cuda"
torch.manual_seed(0)
lib = native.cuda_lib()

# ---- K1 single GPU + LIE moments
D = 98304
rows = torch.randn(16, D, device=dev)
w = torch.full((16,), 1 / 16, device=dev)
mem = (torch.rand(2, 16) < 0.6).float(); mem[:, :2] = 1
flat.fedavg(rows, w, bcast=rows.clone(), member=mem, z=torch.tensor([0.74, 0.74]), ref=rows[0].clone(), want_lie=True)

# ---- K1-fused with world = 1 (the protocol runs against this rank's own flags)
Df = 4 * 1024 * 1024
r1 = torch.randn(2, Df, device=dev); part = torch.zeros(1, Df, device=dev); glob = torch.zeros(Df, device=dev)
sig = torch.zeros(2048, dtype=torch.int32, device=dev); ctrl = torch.zeros(128, dtype=torch.int32, device=dev)
PA = ctypes.c_void_p * 1
check(lib, lib.afl_fedavg_fused(ptr(r1), ptr(torch.full((2,), 0.5, device=dev)), ctypes.c_int(2), ctypes.c_int64(Df), ctypes.c_int(1),
                                ctypes.c_int(0), PA(part.data_ptr()), PA(glob.data_ptr()), None, PA(sig.data_ptr()),
                                ctypes.c_uint32(1), ctypes.c_int(0), None, None, None, None, ctypes.c_int(0), ptr(r1), ptr(ctrl),
                                ctypes.c_void_p(0), ctypes.c_void_p(0), ctypes.c_void_p(0), stream_ptr(r1)), "afl_fedavg_fused")

# ---- K6
sel = torch.randn(128, D, device=dev)
flat.coord_select(sel, "median")
out = torch.zeros(D, device=dev)
flat.coord_select_ptrs([sel[c].data_ptr() for c in range(128)], 0, D, [out.data_ptr()], "trimmed", 12, dev)

# ---- one Gram for all attackers (config 4 shape: 112 genuine rows, 16 attackers, RNN D)
cen = torch.randn(D, device=dev) * 0.1
gen = cen[None] + 0.02 * torch.randn(112, D, device=dev)
g = torch.Generator().manual_seed(1)
subs = [torch.randperm(112, generator=g)[:56].tolist() for _ in range(16)]
flat.multi_gram_attack([gen[j].data_ptr() for j in range(112)], cen, subs, ["Min-Max", "Min-Sum", "Opt-Fang", "Min-Max"] * 4, D)

# ---- hyper round + emit (config 2 shape: CNNModel target, 8 clients)
target = models.build("CNNModel")
lay = FlatLayout.from_module(target)
fused = FusedHyperNet.from_module(HyperNetwork(target, 8, 8, 100, False, 2), lay, device=dev)
nat = flat.NativeHyper(fused)
crow = torch.stack([lay.flatten(models.build("CNNModel").state_dict()) for _ in range(8)]).to(dev)
nat.round(list(range(8)), [crow[c].data_ptr() for c in range(8)], 1e-3, 1.0)
nat.emit(list(range(8)))

# ---- ScionFL
flat.scion_cosine_distances(torch.randn(64, D, device=dev), torch.ones(D, device=dev), seed=3)

# ---- slice-owned ScionFL scoring over row pointers, per-round indices, single-block ROC-AUC
rs = torch.randn(64, D, device=dev)
bits, counts, minmax = flat.scion_quantize(rs, torch.ones(D, device=dev), 3, row0=0)
lo, hi = flat.scion_clipped_levels(counts, minmax, float(D))
flat.scion_scores_ptrs([bits[i].data_ptr() for i in range(64)], D // 32, 0, D // 32, torch.ones(D, device=dev), torch.stack([lo, hi], 1))
flat.round_indices(16384, torch.full((8,), 14000, dtype=torch.int32, device=dev), torch.arange(8, dtype=torch.int32, device=dev), 5, 14000, 1)
flat.roc_auc(torch.rand(4096, device=dev), (torch.rand(4096, device=dev) < 0.3).float())

# ---- ICU training (interpreter + tensor-core trainer) and inference
name = "TransformerModel"
layT = FlatLayout.from_module(models.build(name))
C, n = 8, 640
rowsT = torch.stack([layT.flatten(models.build(name).state_dict()) for _ in range(C)]).to(dev)
data = synthetic_icu(8192, 1).packed().to(dev)
idx = make_round_indices(8192, [n] * C, 1, torch.Generator(device=dev).manual_seed(0), dev)
nd = torch.full((C,), n, dtype=torch.int32, device=dev); cid = torch.arange(C, dtype=torch.int32, device=dev)
prog = build_program(name, layT, S=8, train=True, dropout=True).to(dev)
plan = T.make_plan(prog, 128, C, layT.padded, dev)
m, v = torch.zeros_like(rowsT), torch.zeros_like(rowsT)
T.train_round(prog, plan, rowsT.clone(), m, v, torch.zeros(plan.chunk * plan.G, layT.padded, device=dev), data, idx, nd, cid, 128, 1)
m.zero_(); v.zero_()
TC.train_round_tc(layT, rowsT.clone(), m, v, torch.zeros_like(rowsT), torch.zeros(C, TC.stash_floats_per_client(dev), device=dev),
                  data, idx, nd, cid, 128, 1)
iprog = build_program(name, layT, S=16, train=False).to(dev)
T.infer(iprog, rowsT[:1].contiguous(), data[:4096].contiguous())

# ---- one HAR optimiser step
from attackfl_b200.ops.har import HarNativeTrainer
mod = models.TransformerClassifier()
layH = FlatLayout.from_module(mod)
rowH = layH.flatten(mod.state_dict(), device=dev)
tr = HarNativeTrainer(layH, dev, max_batch=128)
x, y = torch.randn(128, 561, device=dev), torch.randint(0, 6, (128,), device=dev)
grad = torch.zeros_like(rowH)
tr.grad_step(rowH, grad, x, y, seed=1, client=0, step=1, dropout=True)
flat.adam_flat(rowH, grad, torch.zeros_like(rowH), torch.zeros_like(rowH), 1e-3, 1)
torch.cuda.synchronize()
print("ok", native.launch_count(), "launches")
