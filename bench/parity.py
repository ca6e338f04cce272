#!/usr/bin/env python
"""Post-attack accuracy parity: ROC-AUC trajectories of this framework vs the UNMODIFIED reference (driven through the
in-memory pika stand-in, ``baseline/harness_stub.py``) for every attack, over several seeds, same synthetic ICU data.

    python bench/parity.py [--seeds 5] [--rounds 6] [--clients 6] [--out profiles/parity_r2.json]

Per attack in {none, Random, LIE, Min-Max, Min-Sum, Opt-Fang} and per seed both arms run the same configuration (one
attacker = the last client, attacking from round 2, genuine-rate 0.5, TransformerModel, fedavg).  Our arm runs with ``server.common-init: false`` (every client starts from its own random initialisation, like the
reference; with the default common init the clean ROC-AUC settles ~0.01 lower because averaging independent
initialisations shrinks the first global model, which regularises the later rounds).  Dropout and the
model initialisations differ between the arms by construction (the reference seeds nothing on the clients), so parity
is statistical: mean +- 95 % CI of the per-round ROC-AUC (the reference's only metric, src/Validation.py:119-120).
Runs on the CPU (reference: torch CPU; ours: the CPU twin of the kernels) so it needs no GPU time.
"""
import argparse
import contextlib
import io
import json
import math
import os
import sys
import tempfile

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

ATTACKS = [("none", None, None), ("Random", "Random", [0.5]), ("LIE", "LIE", [0.74]), ("Min-Max", "Min-Max", None),
           ("Min-Sum", "Min-Sum", None), ("Opt-Fang", "Opt-Fang", None)]


def ours(attack, args, seed, a):
    import torch

    from attackfl_b200 import from_dict
    from attackfl_b200.data import synthetic_icu
    from attackfl_b200.engine import ClientSpec, Federation
    from attackfl_b200.log import set_quiet
    set_quiet(True)
    torch.manual_seed(seed)
    cfg = from_dict({"server": {"num-round": a.rounds, "clients": a.clients, "mode": "fedavg", "model": "TransformerModel",
                                "data-name": "ICU", "genuine-rate": 0.5, "random-seed": seed, "common-init": a.common_init,
                                "data-distribution": {"num-data-range": [a.num_data, a.num_data + 100]}},
                     "learning": {"epoch": a.epochs, "batch-size": 64, "learning-rate": 0.004, "clip-grad-norm": 1.0},
                     "runtime": {"synthetic_train_rows": a.train_rows, "synthetic_test_rows": a.test_rows}})
    specs = [ClientSpec(i) for i in range(a.clients)]
    if attack:
        specs[-1] = ClientSpec(a.clients - 1, True, attack, 2, args)
    with contextlib.redirect_stdout(io.StringIO()):
        fed = Federation(cfg, specs, device="cpu", train_set=synthetic_icu(a.train_rows, 1), test_set=synthetic_icu(a.test_rows, 102),
                         workdir=tempfile.mkdtemp(prefix="afl_par_"), save=False, verbose=False)
        res = fed.run()
    return [r.metric for r in res if r.ok]


def reference(attack, args, seed, a):
    import random

    import torch

    from baseline import harness_stub
    torch.manual_seed(seed)
    random.seed(seed)
    out = harness_stub.run("fedavg", "TransformerModel", a.clients, a.rounds, attack, 2, args,
                           (a.num_data, a.num_data + 100), a.epochs, 64, a.train_rows, a.test_rows)
    return out["roc_auc"]


def stats(runs, rounds):
    out = []
    for r in range(rounds):
        xs = [run[r] for run in runs if len(run) > r and run[r] == run[r]]
        if not xs:
            out.append(None); continue
        m = sum(xs) / len(xs)
        sd = math.sqrt(sum((x - m) ** 2 for x in xs) / max(len(xs) - 1, 1))
        out.append({"mean": round(m, 4), "ci95": round(1.96 * sd / math.sqrt(len(xs)), 4), "n": len(xs)})
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--seeds", type=int, default=5); ap.add_argument("--rounds", type=int, default=6)
    ap.add_argument("--clients", type=int, default=6); ap.add_argument("--epochs", type=int, default=2)
    ap.add_argument("--num-data", dest="num_data", type=int, default=500)
    ap.add_argument("--train-rows", dest="train_rows", type=int, default=2048)
    ap.add_argument("--test-rows", dest="test_rows", type=int, default=1024)
    ap.add_argument("--common-init", dest="common_init", action="store_true",
                    help="one common round-1 initialisation (this framework's default) instead of the reference's independent client inits")
    ap.add_argument("--out", default="profiles/parity_r2.json")
    a = ap.parse_args()
    table = {"config": vars(a), "attacks": {}}
    for name, attack, args in ATTACKS:
        rec = {"ours": [], "reference": []}
        for seed in range(1, a.seeds + 1):
            rec["ours"].append(ours(attack, args, seed, a))
            try:
                rec["reference"].append(reference(attack, args, seed, a))
            except Exception as e:                               # noqa: BLE001  (e.g. the reference's LIE k=1 NaN)
                rec["reference"].append([])
                rec.setdefault("reference_errors", []).append(repr(e)[:200])
            print(name, seed, "ours", [round(x, 3) for x in rec["ours"][-1]], "ref", [round(x, 3) for x in rec["reference"][-1]], flush=True)
        rec["ours_stats"] = stats(rec["ours"], a.rounds)
        rec["reference_stats"] = stats(rec["reference"], a.rounds)
        table["attacks"][name] = rec
        with open(a.out, "w") as fh:
            json.dump(table, fh, indent=1)
    print("wrote", a.out)


if __name__ == "__main__":
    main()
