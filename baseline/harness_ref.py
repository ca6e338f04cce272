"""Reference arm of ``bench.py``: the UNMODIFIED filrg/attackFL code on the same box, same metric, same config.

What runs (all of it the reference's own code, imported from ``baseline/_ref``):
  * rank 0 hosts ``server.Server`` (``on_request`` -> ``process_consumer`` -> ``avg_all_parameters`` ->
    ``Validation.test`` -> ``notify_clients``);
  * every rank hosts its 8 ``RpcClient``s, which receive the pickled START, run ``client.train_ICU`` on the rank's
    GPU through the reference's DataLoader (``genuine_training``) or ``create_LIE_state_dict`` (attackers), and
    publish the pickled UPDATE.
The RabbitMQ broker is replaced by the in-memory ``pika`` stand-in (``baseline/stubs``); between ranks the
message bodies travel with ``torch.distributed`` object collectives (NCCL).  That is generous to the reference: no
broker hop, no 0.5 s client polling (``src/RpcClient.py:46-51``).  The only source edit is the documented one-token
``fedavg`` fix, applied in memory by ``ref_loader``.
"""
from __future__ import annotations

import contextlib
import io
import os
import sys
import tempfile
import time
import uuid


def _write_datasets(workdir: str, train_rows: int, test_rows: int):
    """Synthetic ICU-shaped data pickled as the reference's own ``src.Model.ICUData`` (Model.py:9-24)."""
    import gzip
    import pickle

    import pandas as pd

    from attackfl_b200.data import synthetic_icu
    from baseline import ref_loader
    RefICU = ref_loader.load_model().ICUData
    vc, lc = [f"v{i}" for i in range(7)], [f"l{i}" for i in range(16)]

    def to_ref(ds):
        df = pd.DataFrame(ds.vitals.numpy(), columns=vc)
        for i, c in enumerate(lc):
            df[c] = ds.labs[:, i].numpy()
        df["label"] = ds.labels.numpy()
        return RefICU(df, vc, lc, "label")

    os.makedirs(os.path.join(workdir, "data"), exist_ok=True)
    with gzip.open(os.path.join(workdir, "train_dataset.pkl.gz"), "wb") as fh:
        pickle.dump(to_ref(synthetic_icu(train_rows, 1)), fh)
    with gzip.open(os.path.join(workdir, "data", "test_dataset.pkl.gz"), "wb") as fh:
        pickle.dump(to_ref(synthetic_icu(test_rows, 102)), fh)


def run_bench(args, bench_config, clients_per_gpu: int, train_rows: int, test_rows: int) -> dict:
    """bench.py --impl reference: BASELINE config 3 (one LIE attacker per 8 clients)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    C = clients_per_gpu * world
    attackers = {c: ("LIE", 2, [0.74]) for c in range(C) if c % clients_per_gpu == clients_per_gpu - 1}
    return run_config(bench_config(C), attackers, args.steps, args.warmup, clients_per_gpu, train_rows, test_rows)


def run_config(cfg: dict, attackers: dict, steps: int, warmup_req: int, clients_per_gpu: int, train_rows: int,
               test_rows: int, min_warmup: int = 3) -> dict:
    """Drives the unmodified reference for ``warmup + steps`` rounds of ``cfg`` (any mode / model) and times the
    last ``steps`` rounds.  ``attackers``: client index -> (mode, first attack round, args)."""
    import types
    args = types.SimpleNamespace(steps=steps, warmup=warmup_req)
    bench_config = lambda n: cfg
    import torch
    import torch.distributed as dist

    from baseline import ref_loader
    if not ref_loader.ensure_installed():
        return {"impl": "reference", "unavailable": "baseline/_ref is not installed (run baseline/install_ref.sh)"}
    cpu_mode = os.environ.get("AFL_REF_DEVICE") == "cpu"          # plumbing test of this harness without a GPU
    if not torch.cuda.is_available() and not cpu_mode:
        return {"impl": "reference", "unavailable": "no CUDA device"}
    os.environ.setdefault("TQDM_DISABLE", "1")
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = "cpu" if cpu_mode else f"cuda:{local}"
    if not cpu_mode:
        torch.cuda.set_device(dev)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29512")
        if cpu_mode:
            dist.init_process_group("gloo", rank=rank, world_size=world)
        else:
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=torch.device(dev))
    warmup = max(args.warmup, min_warmup)
    n_rounds = warmup + args.steps
    C = clients_per_gpu * world
    cfg = bench_config(C)
    cfg["server"]["num-round"] = n_rounds
    if cpu_mode:                                                   # keep the CPU plumbing test short
        cfg["server"]["data-distribution"]["num-data-range"] = [200, 300]
        cfg["learning"]["epoch"] = 1
        train_rows, test_rows = 512, 256
    cwd = os.getcwd()
    workdir = tempfile.mkdtemp(prefix=f"afl_ref_r{rank}_")
    _write_datasets(workdir, train_rows, test_rows)
    sink = io.StringIO()
    try:
        with contextlib.redirect_stdout(sink), contextlib.redirect_stderr(io.StringIO()):
            server_mod, client_mod = ref_loader.load_server_client(workdir, cfg, device=dev, patch_fedavg=True)
        import pika
        my = list(range(rank * clients_per_gpu, (rank + 1) * clients_per_gpu))
        ids = [uuid.UUID(int=c + 1) for c in range(C)]
        with contextlib.redirect_stdout(sink):
            clients = {}
            for c in my:
                rc = client_mod.RpcClient(ids[c], "127.0.0.1", "admin", "admin", client_mod.train_on_device, dev)
                if c in attackers:
                    rc.set_attack_config(*attackers[c])
                clients[c] = rc
            server = server_mod.Server() if rank == 0 else None

        def drain_rpc():
            q = pika.QUEUES["rpc_queue"]
            out = list(q)
            q.clear()
            return out

        def to_server(bodies_by_rank):
            """rank 0: feed UPDATE/REGISTER bodies in client order; returns False once the server exits."""
            alive = True
            if rank == 0:
                try:
                    for bodies in bodies_by_rank:
                        for b in bodies:
                            server.on_request(pika.Channel(), pika._Method(0), pika._Props(), b)
                except SystemExit:
                    alive = False
            return alive

        def gather(bodies):
            if world == 1:
                return [bodies]
            out = [None] * world if rank == 0 else None
            dist.gather_object(bodies, out, dst=0)
            return out

        def scatter_starts():
            """START (or STOP) bodies published by the server -> owning ranks."""
            if rank == 0:
                per_rank = [dict() for _ in range(world)]
                for c in range(C):
                    q = pika.QUEUES[f"reply_{ids[c]}"]
                    if q:
                        per_rank[c // clients_per_gpu][c] = q.popleft()
            else:
                per_rank = None
            if world == 1:
                return per_rank[0]
            got = [None]
            dist.scatter_object_list(got, per_rank, src=0)
            return got[0]

        # ---- registration
        with contextlib.redirect_stdout(sink):
            for c in my:
                clients[c].send_to_server({"action": "REGISTER", "client_id": ids[c], "message": "Hello from Client!",
                                           "attack": c in attackers})
            to_server(gather(drain_rpc()))

        def one_round() -> bool:
            with contextlib.redirect_stdout(sink), contextlib.redirect_stderr(io.StringIO()):
                starts = scatter_starts()
                for c in my:
                    if c in starts:
                        clients[c].response_message(starts[c])
                alive = to_server(gather(drain_rpc()))
            if world > 1:
                flag = torch.tensor([1 if alive else 0], device=dev)
                dist.broadcast(flag, src=0)
                alive = bool(flag.item())
            sink.seek(0); sink.truncate(0)
            return alive

        def barrier():
            if not cpu_mode:
                torch.cuda.synchronize()
            if world > 1:
                dist.barrier()

        class _WallEvent:                                          # CPU plumbing mode: wall clock instead of events
            def __init__(self, enable_timing=True): self.t = 0.0
            def record(self): self.t = time.perf_counter()
            def elapsed_time(self, other): return (other.t - self.t) * 1e3
        Event = _WallEvent if cpu_mode else torch.cuda.Event

        for _ in range(warmup):
            one_round()
        barrier()
        ev = []
        t0 = time.perf_counter()
        for _ in range(args.steps):
            e0, e1 = Event(enable_timing=True), Event(enable_timing=True)
            e0.record()
            one_round()
            e1.record()
            ev.append((e0, e1))
        barrier()
        wall = time.perf_counter() - t0
        dev_ms = sum(a.elapsed_time(b) for a, b in ev)
        t = torch.tensor([dev_ms, wall], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dev_ms, wall = float(t[0]), float(t[1])
        auc = None
        if rank == 0:
            try:
                with open(os.path.join(workdir, "app.log")) as fh:
                    aucs = [ln.rsplit("ROC_AUC:", 1)[1].strip() for ln in fh if "ROC_AUC:" in ln]
                auc = float(aucs[-1]) if aucs else None
            except Exception:                                  # noqa: BLE001
                auc = None
    finally:
        os.chdir(cwd)
    out = {
        "metric": "fl_client_rounds_per_sec", "value": round(C * args.steps / (dev_ms / 1e3), 4),
        "unit": "client-rounds/s", "rounds_per_sec": round(args.steps / (dev_ms / 1e3), 5), "n_gpus": world,
        "steps": args.steps, "warmup": warmup, "ms_per_step": round(dev_ms / args.steps, 2), "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "fp32", "data": "synthetic", "impl": "reference",
        "config": {"model": cfg["server"]["model"], "mode": cfg["server"]["mode"],
                   "attack": {str(k): list(v) for k, v in attackers.items()}, "clients": C, "clients_per_gpu": clients_per_gpu, "local_epochs": 5,
                   "batch_size": 128, "num_data_range": [12000, 15000], "parallelism": f"fl-clients x{world}",
                   "transport": "in-memory pika stand-in + torch.distributed object collectives (no broker)",
                   "patch": "server.py:289 if->elif (fedavg crash fix, applied in memory)"},
        "e2e": {"value": round(C * args.steps / wall, 4), "unit": "client-rounds/s",
                "note": "the reference loop already moves every batch host->device and every loss device->host"},
        "gpu_launches": 0, "roc_auc_last": auc,
    }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out if rank == 0 else {}
