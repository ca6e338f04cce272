#!/usr/bin/env python
"""Semantics / accuracy-parity oracle: the UNMODIFIED reference driven in ONE process on CPU.

    python baseline/harness_stub.py --mode fedavg --model TransformerModel --clients 3 --rounds 3 [--attack LIE ...]

``server.py`` + ``src/RpcClient.py`` run over the in-memory ``pika`` stand-in; messages are pumped
by ``ref_loader.pump``.  Prints the per-round ROC_AUC lines the reference logged to ``app.log``.
"""
import argparse
import contextlib
import io
import json
import os
import sys
import tempfile
import uuid

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))


def run(mode="fedavg", model="TransformerModel", clients=3, rounds=2, attack=None, attack_round=2, attack_args=None,
        num_data=(200, 260), epochs=1, batch=64, train_rows=512, test_rows=256, quiet=True):
    from baseline import harness_ref, ref_loader
    os.environ.setdefault("TQDM_DISABLE", "1")
    cfg = {"name": "stub", "server": {"num-round": rounds, "clients": clients, "mode": mode, "model": model, "data-name": "ICU",
                                      "parameters": {"load": False}, "validation": True, "genuine-rate": 0.5, "random-seed": 1,
                                      "data-distribution": {"num-data-range": list(num_data)},
                                      "hyper-detection": {"enable": False, "cosine-search": 10, "n_components": 3,
                                                          "eps": 0.007, "min_samples": 3}},
           "rabbit": {"address": "127.0.0.1", "username": "admin", "password": "admin"}, "log_path": ".",
           "learning": {"epoch": epochs, "learning-rate": 0.004, "hyper-lr": 0.001, "momentum": 0.5, "batch-size": batch,
                        "clip-grad-norm": 1.0}}
    cwd = os.getcwd()
    work = tempfile.mkdtemp(prefix="afl_stub_")
    harness_ref._write_datasets(work, train_rows, test_rows)
    sink = io.StringIO()
    try:
        with (contextlib.redirect_stdout(sink) if quiet else contextlib.nullcontext()), contextlib.redirect_stderr(io.StringIO()):
            server_mod, client_mod = ref_loader.load_server_client(work, cfg, device="cpu")
            server = server_mod.Server()
            cl = []
            for c in range(clients):
                rc = client_mod.RpcClient(uuid.UUID(int=c + 1), "127.0.0.1", "admin", "admin", client_mod.train_on_device, "cpu")
                is_att = attack is not None and c == clients - 1
                if is_att:
                    rc.set_attack_config(attack, attack_round, attack_args)
                rc.send_to_server({"action": "REGISTER", "client_id": rc.client_id, "message": "Hello from Client!", "attack": is_att})
                cl.append(rc)
            ref_loader.pump(server, cl)
        log = open(os.path.join(work, "app.log")).read()
    finally:
        os.chdir(cwd)
    aucs = [float(ln.rsplit("ROC_AUC:", 1)[1]) for ln in log.splitlines() if "ROC_AUC:" in ln]
    return {"roc_auc": aucs, "workdir": work}


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--mode", default="fedavg"); ap.add_argument("--model", default="TransformerModel")
    ap.add_argument("--clients", type=int, default=3); ap.add_argument("--rounds", type=int, default=2)
    ap.add_argument("--attack", default=None); ap.add_argument("--attack_round", type=int, default=2)
    ap.add_argument("--attack_args", type=float, nargs="+", default=None)
    a = ap.parse_args()
    print(json.dumps(run(a.mode, a.model, a.clients, a.rounds, a.attack, a.attack_round, a.attack_args)))
