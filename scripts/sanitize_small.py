"""Small-shape run of every kernel family for compute-sanitizer (memcheck / racecheck / synccheck)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from attackfl_b200 import FlatLayout, models
from attackfl_b200.data import make_round_indices, synthetic_icu
from attackfl_b200.models.hyper import FusedHyperNet
from attackfl_b200.models.programs import build_program
from attackfl_b200.ops import flat, gram_native, train as T
dev = "cuda"
torch.manual_seed(0)
ds = synthetic_icu(512, 1); data = ds.packed().to(dev)
for name, S, G in (("TransformerModel", 8, 2), ("RNNModel", 8, 2), ("CNNModel", 4, 2)):
    lay = FlatLayout.from_module(models.build(name))
    prog = build_program(name, lay, S=S, train=True, dropout=True).to(dev)
    C = 2
    rows = torch.stack([lay.flatten(models.build(name).state_dict()) for _ in range(C)]).to(dev)
    m, v = torch.zeros_like(rows), torch.zeros_like(rows)
    gpart = torch.zeros(C * G, lay.padded, device=dev)
    idx = make_round_indices(512, [40, 33], 1, torch.Generator(device=dev).manual_seed(0), dev)
    T.train_round(prog, T.TrainPlan(G, 16 // (S * G) or 1, S, C), rows, m, v, gpart, data, idx,
                  torch.tensor([40, 33], dtype=torch.int32, device=dev), torch.arange(C, dtype=torch.int32, device=dev), 16, 1, seed=1)
    ip = build_program(name, lay, S=S, train=False).to(dev)
    T.infer(ip, rows, data[:37].contiguous())
    w = torch.full((C,), 0.5, device=dev)
    flat.fedavg(rows, w, bcast=rows, member=torch.ones(1, C), z=torch.tensor([0.74]), ref=rows[0].clone(), want_lie=True)
    flat.coord_select(torch.randn(5, lay.padded, device=dev), "median")
a = torch.randn(18, 4100, device=dev)
gram_native.gram_tcgen05(a, "tf32"); gram_native.gram_tcgen05(a, "bf16")
h = models.HyperNetwork(models.TransformerModel(), 3, 8, 100, False, 2)
nat = flat.NativeHyper(FusedHyperNet.from_module(h, device=dev))
th = nat.emit([0, 1, 2]); nat.step(1, th[0].clone(), 1e-3, 1.0)
flat.craft_random(th[0].clone(), None, 0.5, 1, 2)
torch.cuda.synchronize()
print("sanitize workload ok")
