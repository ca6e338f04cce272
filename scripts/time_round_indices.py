"""GPU time of the per-round index generation (data.make_round_indices: argsort(rand) subsets and per-epoch shuffles)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from attackfl_b200.data import make_round_indices
g = torch.Generator(device="cuda").manual_seed(0)
for C in (8, 32, 128):
    sizes = [12000 + (i * 37) % 3000 for i in range(C)]
    for _ in range(3):
        make_round_indices(16384, sizes, 5, g, "cuda")
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(10):
        make_round_indices(16384, sizes, 5, g, "cuda")
    b.record(); torch.cuda.synchronize()
    print(f"C={C}: make_round_indices {a.elapsed_time(b) / 10:.3f} ms per round")
