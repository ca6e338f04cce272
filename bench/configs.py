#!/usr/bin/env python
"""BASELINE.json configs 2-4 on one node, both arms.

    python bench/configs.py --config 2|3|4 [--attack Min-Max|Min-Sum|Opt-Fang] [--clients N] [--rounds K]
                            [--warmup W] [--impl ours|reference]
    (multi-GPU: python -m torch.distributed.run --nproc-per-node N bench/configs.py ...)

  2: CNNModel, mode hyper, 8 clients, genuine-rate 0.5
  3: TransformerModel, fedavg, 8 clients per GPU, LIE z=0.74          (same as bench.py)
  4: RNNModel, fedavg, 128 clients, Min-Max / Min-Sum / Opt-Fang       (Gram-GEMM path)
  6: HAR TransformerClassifier (T = 561), fedavg, 8 clients, 1 local epoch of 1,024 samples, batch 128 — not a
     BASELINE.json config (the reference's HAR client path raises); `--backend torch` is the PyTorch-module arm
One JSON line: FL rounds/sec, device-timed (CUDA events per round, max over ranks), shipped learning config.
"""
import argparse
import contextlib
import io
import json
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench as headline  # noqa: E402  (repo-root bench.py: shared config + clock sampler)


def make(config: int, clients: int, attack: str):
    cfg = headline.bench_config(clients)
    attackers = {}
    if config == 2:
        cfg["server"].update(mode="hyper", model="CNNModel")
    elif config == 3:
        attackers = {c: ("LIE", 2, [0.74]) for c in range(clients) if c % 8 == 7}
    elif config == 4:
        cfg["server"].update(mode="fedavg", model="RNNModel")
        attackers = {c: (attack, 2, None) for c in range(clients) if c % 8 == 7}
    elif config == 6:
        cfg["server"].update(mode="fedavg", model="TransformerClassifier")
        cfg["server"]["data-name"] = "HAR"
        cfg["server"]["data-distribution"] = {"num-data-range": [1024, 1024]}
        cfg["learning"].update(epoch=1)
        attackers = {c: ("LIE", 2, [0.74]) for c in range(clients) if c % 8 == 7}
    else:
        raise SystemExit("--config must be 2, 3, 4 or 6")
    return cfg, attackers


def run_ours(cfg_d, attackers, rounds, warmup, backend="auto"):
    import torch

    from attackfl_b200 import from_dict
    from attackfl_b200.data import synthetic_icu
    from attackfl_b200.engine import ClientSpec, Federation
    from attackfl_b200.log import set_quiet
    from attackfl_b200.ops import native
    from attackfl_b200.parallel import ctx as pctx
    set_quiet(True)
    ctx = pctx.init_from_env("cuda")
    cfg = from_dict(cfg_d)
    cfg.runtime.backend = backend
    C = cfg.server.clients
    specs = [ClientSpec(i, True, *attackers[i]) if i in attackers else ClientSpec(i) for i in range(C)]
    if cfg.server.data_name == "HAR":
        from attackfl_b200.data import synthetic_har
        train, test = synthetic_har(4096, 1), synthetic_har(1024, 102)
    else:
        train, test = synthetic_icu(headline.TRAIN_ROWS, 1), synthetic_icu(headline.TEST_ROWS, 102)
    fed = Federation(cfg, specs, ctx=ctx, train_set=train, test_set=test, workdir=tempfile.mkdtemp(), save=False,
                     verbose=False)
    if ctx.distributed and cfg.runtime.collectives == "fused":
        from attackfl_b200.parallel import symm
        symm.attach(fed)
    sink = io.StringIO()
    for _ in range(warmup):
        with contextlib.redirect_stdout(sink):
            fed.run_round()
    torch.cuda.synchronize(); ctx.barrier()
    native.reset_launch_count()
    ev, phases, metric, trace = [], {}, None, []
    for _ in range(rounds):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        with contextlib.redirect_stdout(sink):
            r = fed.run_round()
        e1.record()
        ev.append((e0, e1)); metric = r.metric
        trace.append(round(r.metric, 4) if (r.ok and r.metric == r.metric) else None)
        for k, v in r.timings_ms.items():
            phases[k] = phases.get(k, 0.0) + v / rounds
    torch.cuda.synchronize()
    ms = sum(a.elapsed_time(b) for a, b in ev)
    t = torch.tensor([ms], device=ctx.device); ctx.all_reduce_max(t); ms = float(t.item())
    out = {"impl": "ours", "metric": "fl_rounds_per_sec", "value": round(rounds / (ms / 1e3), 4), "ms_per_round": round(ms / rounds, 2),
           "n_gpus": ctx.world, "clients": C, "rounds": rounds, "warmup": warmup, "model": cfg.server.model,
           "mode": cfg.server.mode, "attackers": {str(k): list(v) for k, v in attackers.items()},
           "phases_ms": {k: round(v, 3) for k, v in phases.items()}, "gpu_launches": native.launch_count(),
           "metric_last": metric, "metric_trace": trace, "trainer": getattr(fed, "trainer", None), "plan": ({"G": fed.plan.G, "M": fed.plan.M, "S": fed.plan.S, "chunk": fed.plan.chunk} if getattr(fed, "plan", None)
                                             else {"trainer": "har-native" if fed.har_trainer is not None else "torch"})}
    return out if ctx.rank == 0 else {}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", type=int, required=True)
    ap.add_argument("--attack", default="Min-Max")
    ap.add_argument("--clients", type=int, default=0)
    ap.add_argument("--rounds", type=int, default=3)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours")
    ap.add_argument("--backend", default="auto", help="runtime.backend of our arm: auto | native | torch")
    a = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    clients = a.clients or {2: 8, 3: 8 * world, 4: 128, 6: 8 * world}[a.config]
    cfg, attackers = make(a.config, clients, a.attack)
    if a.impl == "reference":
        from baseline import harness_ref
        out = harness_ref.run_config(cfg, attackers, a.rounds, a.warmup, clients // world, headline.TRAIN_ROWS, headline.TEST_ROWS,
                                     min_warmup=1)
        if out:
            out["rounds_per_sec"] = out.get("rounds_per_sec")
    else:
        out = run_ours(cfg, attackers, a.rounds, a.warmup, a.backend)
        if out:
            out["backend"] = a.backend
    if out:
        out["bench_config"] = a.config
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
