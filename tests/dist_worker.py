"""Worker for the multi-process tests: ``python -m torch.distributed.run ... tests/dist_worker.py OUT MODE [ATTACK]``."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from attackfl_b200 import from_dict  # noqa: E402
from attackfl_b200.engine import ClientSpec, Federation  # noqa: E402
from attackfl_b200.parallel import ctx as pctx  # noqa: E402


def main():
    out, mode = sys.argv[1], sys.argv[2]
    attack = sys.argv[3] if len(sys.argv) > 3 else ""
    device = os.environ.get("AFL_TEST_DEVICE", "cpu")
    ctx = pctx.init_from_env(device)
    n = int(os.environ.get("AFL_TEST_CLIENTS", "6"))
    cfg = from_dict({"server": {"num-round": int(os.environ.get("AFL_TEST_ROUNDS", "2")), "clients": n, "mode": mode, "model": "TransformerModel", "data-name": "ICU",
                                "data-distribution": {"num-data-range": [200, 260]}, "random-seed": 1},
                     "learning": {"epoch": 1, "batch-size": 64, "learning-rate": 0.004, "hyper-lr": 0.001,
                                  "clip-grad-norm": 1.0},
                     "runtime": {"synthetic_train_rows": 512, "synthetic_test_rows": 256,
                                 "collectives": os.environ.get("AFL_TEST_COLLECTIVES", "nccl")}})
    specs = [ClientSpec(i) for i in range(n)]
    if attack:
        specs[n - 1] = ClientSpec(n - 1, True, attack, 2, [0.74] if attack in ("LIE", "Random") else None)
    fed = Federation(cfg, specs, ctx=ctx, workdir=os.path.dirname(out), save=(ctx.rank == 0), verbose=False)
    if ctx.distributed and cfg.runtime.collectives == "fused":
        from attackfl_b200.parallel import symm
        symm.attach(fed)
        if os.environ.get("AFL_TEST_JITTER"):
            # randomised per-rank delays (up to ~2 ms of device time) in front of every fused collective: the in-kernel
            # flag barriers must not depend on the ranks arriving together
            import random
            rng = random.Random(1234 + 17 * ctx.rank)
            comm = fed.fused_comm
            for name in ("fedavg", "gram_attack"):
                inner = getattr(comm, name)

                def delayed(*a, _inner=inner, **k):
                    torch.cuda._sleep(rng.randrange(0, 4_000_000))
                    return _inner(*a, **k)
                setattr(comm, name, delayed)
    res = fed.run()
    head_rows = None
    if mode == "hyper":
        sd = fed.state_dict()                 # a collective: gathers the row shards of a sharded head
        # the head bias in arena order, rebuilt from the exported per-tensor entries (full length on every rank)
        glob = torch.zeros(fed.layout.padded)
        for e in fed.layout.entries:
            glob[e.offset:e.offset + e.numel] = sd["hyper_layers." + e.name.replace(".", "__") + ".bias"].float().cpu()
        head_rows = int(fed.hyper.head_w.shape[0])        # rows of W_h this rank actually stores
    else:
        glob = fed.global_row
    rec = {"rank": ctx.rank, "world": ctx.world, "ok": [r.ok for r in res], "metric": [r.metric for r in res],
           "attackers": [r.attackers for r in res], "checksum": float(glob.double().sum()),
           "absmax": float(glob.abs().max()), "local": fed.local, "fused": fed.fused_comm is not None,
           "kept": [r.info.get("scionfl_kept") for r in res], "head_rows": head_rows, "d_pad": int(fed.layout.padded)}
    with open(f"{out}.rank{ctx.rank}.json", "w") as fh:
        json.dump(rec, fh)
    torch.save(glob.detach().float().cpu(), f"{out}.rank{ctx.rank}.pt")           # full vector: compared element-wise
    fed.close()
    if ctx.distributed:
        torch.distributed.barrier()
        torch.distributed.destroy_process_group()


if __name__ == "__main__":
    main()
