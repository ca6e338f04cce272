"""Short single-launch workload for ncu: one persistent training launch (8 clients x 10 optimiser steps),
one FedAvg and one inference launch.  Usage: python scripts/prof_train.py [Model] [n_data] [epochs]"""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from attackfl_b200 import FlatLayout, models
from attackfl_b200.data import make_round_indices, synthetic_icu
from attackfl_b200.models.programs import build_program
from attackfl_b200.ops import train as T, flat

name = sys.argv[1] if len(sys.argv) > 1 else "TransformerModel"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 1280
epochs = int(sys.argv[3]) if len(sys.argv) > 3 else 1
C = int(os.environ.get("PROF_CLIENTS", "8"))
dev = "cuda"
torch.manual_seed(0)
lay = FlatLayout.from_module(models.build(name))
S_env = int(os.environ.get("PROF_S", "0"))
prog = build_program(name, lay, S=S_env or None, train=True, dropout=True, chain=bool(int(os.environ.get("PROF_CHAIN", "0")))).to(dev)
plan = T.make_plan(prog, 128, C, lay.padded, dev, int(os.environ.get("PROF_G", "0")))
print("plan", plan, "pool KB", prog.pool_floats * 4 / 1024)
rows = torch.stack([lay.flatten(models.build(name).state_dict()) for _ in range(C)]).to(dev)
m, v = torch.zeros_like(rows), torch.zeros_like(rows)
gpart = torch.zeros(plan.chunk * plan.G, lay.padded, device=dev)
data = synthetic_icu(16384, 1).packed().to(dev)
g = torch.Generator(device=dev).manual_seed(0)
idx = make_round_indices(16384, [n] * C, epochs, g, dev)
nd = torch.full((C,), n, dtype=torch.int32, device=dev)
cid = torch.arange(C, dtype=torch.int32, device=dev)
for it in range(3):
    m.zero_(); v.zero_()
    torch.cuda.synchronize(); e0 = torch.cuda.Event(enable_timing=True); e1 = torch.cuda.Event(enable_timing=True)
    e0.record()
    loss, steps = T.train_round(prog, plan, rows, m, v, gpart, data, idx, nd, cid, 128, epochs, seed=it, dropout=True)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    st = int(steps[0])
    print(f"train launch {ms:.3f} ms, {st} steps/client, {1e3 * ms / max(st,1):.1f} us/step, loss {float(loss[0,0])/n/epochs:.4f}")
w = torch.full((C,), 1.0 / C, device=dev)
out, _, _, _ = flat.fedavg(rows, w, bcast=rows)
iprog = build_program(name, lay, S=16 if name != "CNNModel" else 4, train=False).to(dev)
probs = T.infer(iprog, out[None].contiguous(), data[:4096].contiguous())
torch.cuda.synchronize()
print("ok", float(probs.mean()))
