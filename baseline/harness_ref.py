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


def _client_process(conn, workdir: str, cfg: dict, dev: str, c: int, attack_cfg):
    """One OS process per client, as the reference is deployed (client.py:134-142): the unmodified ``RpcClient`` handles
    START bodies received from the parent and returns the pickled UPDATE bodies it published."""
    import io as _io
    import contextlib as _ctx
    os.environ.setdefault("TQDM_DISABLE", "1")
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from baseline import ref_loader
    try:
        with _ctx.redirect_stdout(_io.StringIO()), _ctx.redirect_stderr(_io.StringIO()):
            _, client_mod = ref_loader.load_server_client(workdir, cfg, device=dev, patch_fedavg=True)
            import pika
            rc = client_mod.RpcClient(uuid.UUID(int=c + 1), "127.0.0.1", "admin", "admin", client_mod.train_on_device, dev)
            if attack_cfg:
                rc.set_attack_config(*attack_cfg)
        conn.send("ready")
        while True:
            body = conn.recv()
            if body is None:
                break
            with _ctx.redirect_stdout(_io.StringIO()), _ctx.redirect_stderr(_io.StringIO()):
                rc.response_message(body)
            q = pika.QUEUES["rpc_queue"]
            out = list(q)
            q.clear()
            conn.send(out)
    except Exception as e:                                        # noqa: BLE001
        try:
            conn.send(("error", repr(e)))
        except Exception:                                         # noqa: BLE001
            pass


def run_concurrent(cfg: dict, attackers: dict, steps: int, warmup: int, clients_per_gpu: int, train_rows: int,
                   test_rows: int) -> dict:
    """Second reference arm (single node, single GPU): the same unmodified code with the clients running CONCURRENTLY, one OS
    process per client sharing the GPU - how the reference is actually run (client.py:134-142) - instead of one after the
    other in the harness process.  Wall-clock timed (the clients' kernels run in other processes)."""
    import multiprocessing as mp
    import pickle

    import torch

    from baseline import ref_loader
    cpu_mode = os.environ.get("AFL_REF_DEVICE") == "cpu"          # plumbing test without a GPU
    if not ref_loader.ensure_installed() or not (torch.cuda.is_available() or cpu_mode):
        return {"unavailable": "reference or CUDA device missing"}
    os.environ.setdefault("TQDM_DISABLE", "1")
    local = int(os.environ.get("LOCAL_RANK", "0"))
    dev = "cpu" if cpu_mode else f"cuda:{local}"
    sync = (lambda: None) if cpu_mode else torch.cuda.synchronize
    C = clients_per_gpu
    cfg = dict(cfg)
    cfg["server"] = dict(cfg["server"], **{"num-round": warmup + steps, "clients": C})
    cwd = os.getcwd()
    workdir = tempfile.mkdtemp(prefix="afl_refc_")
    _write_datasets(workdir, train_rows, test_rows)
    ctx = mp.get_context("spawn")
    procs, conns = [], []
    sink = io.StringIO()
    try:
        with contextlib.redirect_stdout(sink), contextlib.redirect_stderr(io.StringIO()):
            server_mod, _ = ref_loader.load_server_client(workdir, cfg, device=dev, patch_fedavg=True)
        import pika
        for c in range(C):
            a, b = ctx.Pipe()
            p = ctx.Process(target=_client_process, args=(b, workdir, cfg, dev, c, attackers.get(c)), daemon=True)
            p.start()
            procs.append(p); conns.append(a)
        for a in conns:
            msg = a.recv()
            if msg != "ready":
                return {"unavailable": f"client process failed: {msg}"}
        ids = [uuid.UUID(int=c + 1) for c in range(C)]
        with contextlib.redirect_stdout(sink):
            server = server_mod.Server()
            for c in range(C):
                body = pickle.dumps({"action": "REGISTER", "client_id": ids[c], "message": "Hello from Client!",
                                     "attack": c in attackers})
                server.on_request(pika.Channel(), pika._Method(0), pika._Props(), body)

        def one_round():
            with contextlib.redirect_stdout(sink), contextlib.redirect_stderr(io.StringIO()):
                active = []
                for c in range(C):
                    q = pika.QUEUES[f"reply_{ids[c]}"]
                    if q:
                        conns[c].send(q.popleft())               # all clients start at once
                        active.append(c)
                try:
                    for c in active:
                        got = conns[c].recv()
                        if isinstance(got, tuple):
                            raise RuntimeError(got[1])
                        for b in got:
                            server.on_request(pika.Channel(), pika._Method(0), pika._Props(), b)
                except SystemExit:
                    pass
            sink.seek(0); sink.truncate(0)

        for _ in range(warmup):
            one_round()
        sync()
        t0 = time.perf_counter()
        for _ in range(steps):
            one_round()
        sync()
        wall = time.perf_counter() - t0
    finally:
        os.chdir(cwd)
        for a in conns:
            try:
                a.send(None)
            except Exception:                                     # noqa: BLE001
                pass
        for p in procs:
            p.join(timeout=10)
            if p.is_alive():
                p.terminate()
    return {"value": round(C * steps / wall, 4), "unit": "client-rounds/s", "ms_per_step": round(wall / steps * 1e3, 2),
            "steps": steps, "warmup": warmup, "clients": C, "timing": "wall clock (client kernels run in other processes)",
            "what": "same unmodified reference code, one OS process per client sharing the GPU (client.py:134-142)"}


def run_bench(args, bench_config, clients_per_gpu: int, train_rows: int, test_rows: int, headline_config=None) -> dict:
    """bench.py --impl reference: BASELINE config 3 (one LIE attacker per 8 clients)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    C = clients_per_gpu * world
    attackers = {c: ("LIE", 2, [0.74]) for c in range(C) if c % clients_per_gpu == clients_per_gpu - 1}
    out = run_config(bench_config(C), attackers, args.steps, args.warmup, clients_per_gpu, train_rows, test_rows)
    if out and headline_config is not None and "config" in out:
        # the same ``config`` object as our arm (keys, values, types); arm-specific facts live under impl_detail
        out["impl_detail"] = {k: out["config"][k] for k in ("transport", "patch") if k in out["config"]}
        out["impl_detail"]["attackers"] = out["config"].get("attack")
        out["config"] = headline_config(C, world)
    if out and world == 1 and os.environ.get("AFL_REF_CONCURRENT", "1") != "0" and os.environ.get("AFL_REF_DEVICE") != "cpu":
        # second arm: the reference's clients as concurrent OS processes (its real deployment), fewer rounds
        try:
            out["concurrent_clients"] = run_concurrent(bench_config(C), attackers, min(args.steps, 4), 2, clients_per_gpu,
                                                       train_rows, test_rows)
        except Exception as e:                                    # noqa: BLE001
            out["concurrent_clients"] = {"unavailable": repr(e)}
    return out


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
        auc, auc_trace = None, []
        if rank == 0:
            try:
                with open(os.path.join(workdir, "app.log")) as fh:
                    aucs = [ln.rsplit("ROC_AUC:", 1)[1].strip() for ln in fh if "ROC_AUC:" in ln]
                auc = float(aucs[-1]) if aucs else None
                auc_trace = [float(x) for x in aucs]
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
                # what the stock loop moves per FL round (client.py:88-100, src/Validation.py:98-112), computed from the
                # tensors it copies: every batch of every epoch host->device as (vitals[7], labs[16], label) fp32, the test
                # set once per round; device->host the loss (4 B) and the NaN test (1 B) of every step, the test outputs
                "h2d_bytes_per_step": int(C * cfg["learning"]["epoch"] * sum(cfg["server"]["data-distribution"]["num-data-range"]) / 2 * 24 * 4
                                          + test_rows * 23 * 4),
                "d2h_bytes_per_step": int(C * cfg["learning"]["epoch"] * -(-sum(cfg["server"]["data-distribution"]["num-data-range"]) // (2 * cfg["learning"]["batch-size"])) * 5
                                          + test_rows * 4),
                "note": "the reference loop already moves every batch host->device and every loss device->host; bytes are "
                        "computed from the mean num_data of the configured range"},
        "gpu_launches": 0, "roc_auc_last": auc, "roc_auc_trace": auc_trace,
    }
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return out if rank == 0 else {}
