#!/usr/bin/env python
"""HAR optimiser step (batch 128 x 561 tokens): native trainer vs the PyTorch module (fp32 and TF32 matmul), CUDA
events, per-kernel launch count."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from attackfl_b200 import FlatLayout, models  # noqa: E402
from attackfl_b200.ops import flat, native  # noqa: E402
from attackfl_b200.ops.har import HarNativeTrainer  # noqa: E402


def timeit(fn, iters=10, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(); fn(); b.record()
        torch.cuda.synchronize()
        ts.append(a.elapsed_time(b))
    ts.sort()
    return ts[len(ts) // 2]


def main():
    B = int(os.environ.get("HAR_B", "128"))
    torch.manual_seed(0)
    mod = models.TransformerClassifier().cuda().train()
    lay = FlatLayout.from_module(mod)
    row = lay.flatten(mod.state_dict(), device="cuda")
    tr = HarNativeTrainer(lay, "cuda", max_batch=B)
    x = torch.randn(B, 561, device="cuda"); y = torch.randint(0, 6, (B,), device="cuda")
    grad, m, v = torch.zeros_like(row), torch.zeros_like(row), torch.zeros_like(row)
    state = {"step": 0}

    def native_step():
        state["step"] += 1
        tr.grad_step(row, grad, x, y, seed=1, client=0, step=state["step"], dropout=True)
        flat.adam_flat(row, grad, m, v, 1e-3, state["step"])

    native.reset_launch_count()
    native_step()
    torch.cuda.synchronize()
    launches = native.launch_count()
    out = {"batch": B, "tokens": B * 561, "native_ms": round(timeit(native_step), 3), "native_launches_per_step": launches}
    opt = torch.optim.Adam(mod.parameters(), lr=1e-3)

    def torch_step():
        opt.zero_grad(set_to_none=True)
        torch.nn.functional.cross_entropy(mod(x[:, None, :]), y).backward()
        opt.step()

    for tf32 in (False, True):
        torch.backends.cuda.matmul.allow_tf32 = tf32
        out["torch_tf32_ms" if tf32 else "torch_fp32_ms"] = round(timeit(torch_step), 3)
    torch.backends.cuda.matmul.allow_tf32 = False
    # per-phase breakdown of the native step
    w, g = lay.views(row), lay.views(grad)
    out["native_fwd_ms"] = round(timeit(lambda: tr.forward(w, x, 1, 0, 1, True)), 3)
    # Is the 61-launch Python loop the limit?  Enqueue 16 steps back to back without synchronising: the host time to
    # issue them against the device time to run them (host < device: the GPU never waits for a launch, a CUDA graph of
    # the step would not shorten the round).
    import time
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    t0 = time.perf_counter()
    a.record()
    for _ in range(16):
        native_step()
    b.record()
    host = (time.perf_counter() - t0) * 1e3
    torch.cuda.synchronize()
    out["host_enqueue_ms_per_step"] = round(host / 16, 3)
    out["device_ms_per_step_back_to_back"] = round(a.elapsed_time(b) / 16, 3)
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
