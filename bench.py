#!/usr/bin/env python
"""Headline benchmark: FL rounds/sec (device-timed, max over ranks) — BASELINE.json config 3.

    python bench.py --gpus N --steps K --warmup W [--impl ours|reference]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N ... bench.py --gpus N ...

Config (weak scaling, 8 clients per GPU => 64 clients on 8xB200): ICU ``TransformerModel``, ``mode: fedavg``,
one LIE attacker (``--attack_args 0.74``, attacking from round 2) per 8 clients, the reference's shipped
learning config (5 local epochs, batch 128, Adam lr 0.004, ``num-data-range`` [12000, 15000], server-side
validation every round), fp32 like the reference, synthetic ICU-shaped data and random-init weights.
One "step" = one complete FL round: START fan-out, local training of every client (~525 optimiser steps
each), attack crafting, sample-weighted FedAvg, ROC-AUC validation.  ``value`` = client-rounds per second
summed over all GPUs (``rounds_per_sec`` is reported next to it).

Timing: W >= 3 warm-up rounds, every timed round bracketed by CUDA events on the launching stream, an
L2 flush (256 MiB write) between timed rounds outside the event pairs, max over ranks.  ``e2e`` repeats the
measurement through the public API (`Federation.run_round`) with the round's training pool copied from
pinned host memory every round and the per-client results / ROC-AUC read back to the host, wall-clock
bracketed by barrier + synchronize.
"""
from __future__ import annotations

import argparse
import json
import os
import statistics
import subprocess
import sys
import tempfile
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

CLIENTS_PER_GPU = 8
MODEL = "TransformerModel"
TRAIN_ROWS = 16384
TEST_ROWS = 4096


def bench_config(n_clients: int) -> dict:
    return {
        "name": "bench: BASELINE config 3",
        "server": {"num-round": 10 ** 6, "clients": n_clients, "mode": "fedavg", "model": MODEL, "data-name": "ICU",
                   "validation": True, "genuine-rate": 0.5, "random-seed": 1,
                   "common-init": False,      # reference semantics: every client starts from its own random init

                   "data-distribution": {"num-data-range": [12000, 15000]},
                   "hyper-detection": {"enable": False, "cosine-search": 10, "n_components": 3, "eps": 0.007,
                                       "min_samples": 3},
                   "parameters": {"load": False}},
        "learning": {"epoch": 5, "learning-rate": 0.004, "hyper-lr": 0.001, "momentum": 0.5, "batch-size": 128,
                     "clip-grad-norm": 1.0},
        "rabbit": {"address": "127.0.0.1", "username": "admin", "password": "admin"},
        "log_path": ".",
    }


def headline_config(n_clients: int, n_gpus: int) -> dict:
    """The ``config`` object of the JSON line: identical keys, values and types in BOTH arms (our engine and the
    unmodified reference); everything that is specific to one implementation goes under ``impl_detail``."""
    return {"model": MODEL, "mode": "fedavg", "attack": "LIE z=0.74, 1 attacker per 8 clients from round 2",
            "clients": n_clients, "clients_per_gpu": CLIENTS_PER_GPU, "global_batch": 128 * n_clients, "local_epochs": 5,
            "batch_size": 128, "num_data_range": [12000, 15000], "optimizer_steps_per_client_round": "~525",
            "seq_len": 1, "validation": "ROC-AUC every round", "parallelism": f"fl-clients x{n_gpus} (dp{n_gpus})"}


def _pct(xs, q):
    xs = sorted(xs)
    if not xs:
        return None
    k = min(len(xs) - 1, max(0, int(round(q * (len(xs) - 1)))))
    return round(xs[k], 3)


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index: int = 0):
        self.idx = gpu_index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(self.idx)], stdout=subprocess.PIPE,
                                         stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except Exception:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2]))
            except ValueError:
                continue
            for nm, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        return {"sm_mhz": statistics.median(sm) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def run_ours(args) -> dict:
    import torch
    import torch.distributed as dist

    from attackfl_b200 import from_dict
    from attackfl_b200.data import synthetic_icu
    from attackfl_b200.engine import ClientSpec, Federation
    from attackfl_b200.log import set_quiet
    from attackfl_b200.ops import native
    from attackfl_b200.parallel import ctx as pctx

    set_quiet(True)
    ctx = pctx.init_from_env("cuda")
    assert ctx.device.type == "cuda", "bench.py needs a GPU (use the CPU tests for plumbing)"
    n_gpus = ctx.world
    C = CLIENTS_PER_GPU * n_gpus
    cfg = from_dict(bench_config(C))
    cfg.runtime.synthetic_train_rows, cfg.runtime.synthetic_test_rows = TRAIN_ROWS, TEST_ROWS
    if os.environ.get("ATTACKFL_COLLECTIVES"):
        cfg.runtime.collectives = os.environ["ATTACKFL_COLLECTIVES"]
    specs = [ClientSpec(i, attack=(i % CLIENTS_PER_GPU == CLIENTS_PER_GPU - 1), attack_mode="LIE", attack_round=2,
                        attack_args=[0.74]) if i % CLIENTS_PER_GPU == CLIENTS_PER_GPU - 1 else ClientSpec(i)
             for i in range(C)]
    train = synthetic_icu(TRAIN_ROWS, 1)
    test = synthetic_icu(TEST_ROWS, 102)
    import contextlib
    import io
    workdir = tempfile.mkdtemp(prefix="afl_bench_")
    fed = Federation(cfg, specs, ctx=ctx, train_set=train, test_set=test, workdir=workdir, save=False, verbose=False)
    if ctx.distributed and cfg.runtime.collectives == "fused":
        try:
            from attackfl_b200.parallel import symm
            symm.attach(fed)
        except Exception as e:                                     # noqa: BLE001
            if ctx.rank == 0:
                print(f"[bench] fused collectives unavailable ({e!r}); using NCCL path", file=sys.stderr)
    flush = torch.empty(64 * 1024 * 1024, dtype=torch.float32, device=ctx.device)      # 256 MiB > 126 MB L2
    sink = io.StringIO()

    def one_round():
        with contextlib.redirect_stdout(sink):
            r = fed.run_round()
        return r

    for _ in range(max(args.warmup, 3)):
        one_round()
    # ---------------- device-timed rounds
    torch.cuda.synchronize()
    ctx.barrier()
    sampler = ClockSampler(ctx.local_rank)
    if ctx.rank == 0:
        sampler.start()
    native.reset_launch_count()
    ev = []
    metrics = []
    phases = {}
    agg_ms = []
    for _ in range(args.steps):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        r = one_round()
        e1.record()
        ev.append((e0, e1))
        metrics.append(r.metric)
        for k, v in r.timings_ms.items():
            phases[k] = phases.get(k, 0.0) + v / args.steps
        agg_ms.append(float(r.timings_ms.get("aggregate", 0.0)))
        assert r.ok, "a timed round failed"
    torch.cuda.synchronize()
    launches = native.launch_count()
    per_round = torch.tensor([a.elapsed_time(b) for a, b in ev], device=ctx.device)
    ctx.all_reduce_max(per_round)                                  # per-round max over ranks
    agg_t = torch.tensor(agg_ms, device=ctx.device)
    ctx.all_reduce_max(agg_t)
    per_round, agg_ms = per_round.tolist(), agg_t.tolist()
    dev_ms = float(sum(per_round))
    # synchronous FL: a round lasts as long as its slowest client; with more clients in the job the expected maximum of
    # the per-client num_data (drawn from num-data-range every round) grows, which shows here as the spread between ranks
    tr = torch.tensor([phases.get("train", 0.0), -phases.get("train", 0.0)], device=ctx.device)
    ctx.all_reduce_max(tr)
    train_max, train_min = float(tr[0]), -float(tr[1])
    host_trace = r.info.get("host_trace_ms")
    # ---------------- end-to-end rounds (pinned host -> device every round, results read back)
    pinned = train.packed().pin_memory()
    fed.set_host_train(pinned)
    one_round()
    torch.cuda.synchronize()
    ctx.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        r = one_round()
        _ = (r.metric, r.losses)                                    # already host values (D2H inside run_round)
    torch.cuda.synchronize()
    ctx.barrier()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], device=ctx.device)
    ctx.all_reduce_max(t)
    e2e_s = float(t.item())
    clocks = sampler.stop() if ctx.rank == 0 else {}
    ms_per_step = dev_ms / args.steps
    value = C * args.steps / (dev_ms / 1e3)
    out = {
        "metric": "fl_client_rounds_per_sec", "value": round(value, 3), "unit": "client-rounds/s",
        "rounds_per_sec": round(args.steps / (dev_ms / 1e3), 4),
        "n_gpus": n_gpus, "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": round(ms_per_step, 3),
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic",
        "impl": "ours",
        "config": headline_config(C, n_gpus),
        "impl_detail": {"l2": "flushed with a 256 MiB write between timed rounds", "trainer": fed.trainer,
                        "blocks_per_client": fed.plan.G, "tile_samples": fed.plan.S,
                        "collectives": cfg.runtime.collectives if ctx.distributed else "single-gpu",
                        "fused_attached": fed.fused_comm is not None,
                        "nvls": bool(getattr(fed.fused_comm, "multicast", False)) if fed.fused_comm is not None else False,
                        "launches_per_round": round(launches / max(args.steps, 1), 2),
                        "train_ms_slowest_rank": round(train_max, 3), "train_ms_fastest_rank": round(train_min, 3)},
        "round_ms": {"median": _pct(per_round, 0.5), "p95": _pct(per_round, 0.95), "max": _pct(per_round, 1.0)},
        "aggregate_ms": {"median": _pct(agg_ms, 0.5), "p95": _pct(agg_ms, 0.95), "max": _pct(agg_ms, 1.0)},
        "clocks": clocks,
        "e2e": {"value": round(C * args.steps / e2e_s, 3), "unit": "client-rounds/s",
                "rounds_per_sec": round(args.steps / e2e_s, 4),
                "h2d_bytes_per_step": int(pinned.numel() * 4),
                "d2h_bytes_per_step": int(C * 3 * 4 + 8)},
        "gpu_launches": int(launches),
        "phases_ms": {k: round(v, 3) for k, v in phases.items()},
        "roc_auc_last": metrics[-1],
    }
    if host_trace:
        out["host_trace_ms_last_round"] = host_trace
    fed.close()
    if ctx.distributed:
        dist.barrier()
        dist.destroy_process_group()
    return out if ctx.rank == 0 else {}


def run_reference(args) -> dict:
    try:
        from baseline import harness_ref
    except Exception as e:                                          # noqa: BLE001
        return {"impl": "reference", "unavailable": f"baseline harness import failed: {e!r}"}
    return harness_ref.run_bench(args, bench_config, CLIENTS_PER_GPU, TRAIN_ROWS, TEST_ROWS, headline_config)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", choices=["ours", "reference"], default="ours")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.gpus > 1 and world == 1:
        # convenience: re-launch under torchrun when called directly with --gpus N
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", os.environ.get("MASTER_PORT", "29533"),
               os.path.abspath(__file__), "--gpus", str(args.gpus), "--steps", str(args.steps), "--warmup",
               str(args.warmup), "--impl", args.impl]
        sys.exit(subprocess.call(cmd))
    out = run_reference(args) if args.impl == "reference" else run_ours(args)
    if out:
        print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
