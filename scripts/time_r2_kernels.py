"""CUDA-event timings of the round-2 kernels at BASELINE shapes: hyper round (config 2: CNN target, 8 clients), one-Gram
attack path (config 4: 112 genuine rows, 16 attackers, RNN D)."""
import os, sys, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from attackfl_b200 import FlatLayout, models
from attackfl_b200.models.hyper import FusedHyperNet, HyperNetwork
from attackfl_b200.ops import flat

dev = "cuda"
def timeit(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    return sorted(ts)[len(ts) // 2]

out = {}
for model in ("CNNModel", "TransformerModel"):
    target = models.build(model)
    lay = FlatLayout.from_module(target)
    fused = FusedHyperNet.from_module(HyperNetwork(target, 8, 8, 100, False, 2), lay, device=dev)
    nat = flat.NativeHyper(fused)
    rows = torch.stack([lay.flatten(models.build(model).state_dict()) for _ in range(8)]).to(dev)
    ptrs = [rows[c].data_ptr() for c in range(8)]
    out[f"hyper_round_8clients_{model}_ms"] = round(timeit(lambda: nat.round(list(range(8)), ptrs, 1e-3, 1.0)), 4)
    out[f"hyper_emit_8clients_{model}_ms"] = round(timeit(lambda: nat.emit(list(range(8)))), 4)
    out[f"hyper_step_x8_old_{model}_ms"] = round(timeit(lambda: [nat.step(c, rows[c], 1e-3, 1.0) for c in range(8)]), 4)
    out[f"hyper_bytes_per_client_{model}_MB"] = round(lay.padded * 100 * 4 * 7 / 1e6, 1)
D = 98304
cen = torch.randn(D, device=dev) * 0.1
gen = cen[None] + 0.02 * torch.randn(112, D, device=dev)
g = torch.Generator().manual_seed(1)
subs = [torch.randperm(112, generator=g)[:56].tolist() for _ in range(16)]
modes = ["Min-Max", "Min-Sum", "Opt-Fang", "Min-Max"] * 4
ptrs = [gen[j].data_ptr() for j in range(112)]
out["one_gram_16_attackers_ms"] = round(timeit(lambda: flat.multi_gram_attack(ptrs, cen, subs, modes, D)), 4)
print(json.dumps(out))
