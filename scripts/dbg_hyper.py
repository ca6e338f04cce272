import os, sys, torch
sys.path.insert(0, "/root/repo")
from attackfl_b200 import from_dict
from attackfl_b200.engine import ClientSpec, Federation
from attackfl_b200.parallel import ctx as pctx
mode_coll = sys.argv[1]
ctx = pctx.init_from_env("cuda")
n = 8
cfg = from_dict({"server": {"num-round": 2, "clients": n, "mode": "hyper", "model": "TransformerModel", "data-name": "ICU",
                            "data-distribution": {"num-data-range": [200, 260]}, "random-seed": 1},
                 "learning": {"epoch": 1, "batch-size": 64, "learning-rate": 0.004, "hyper-lr": 0.001, "clip-grad-norm": 1.0},
                 "runtime": {"synthetic_train_rows": 512, "synthetic_test_rows": 256, "collectives": mode_coll}})
fed = Federation(cfg, [ClientSpec(i) for i in range(n)], ctx=ctx, workdir="/tmp", save=False, verbose=False)
if cfg.runtime.collectives == "fused":
    from attackfl_b200.parallel import symm
    symm.attach(fed)
out = {}
for r in range(2):
    res = fed.run_round()
    rows = fed.ctx.all_gather_rows(fed.rows[:fed.CL], fed.C).cpu()
    sd = fed.state_dict()
    out[r] = {"rows": rows, "sd": {k: v.clone() for k, v in sd.items() if not k.startswith("target")}, "metric": res.metric}
if ctx.rank == 0:
    torch.save(out, f"/tmp/dbg_{mode_coll}.pt")
torch.distributed.barrier(); torch.distributed.destroy_process_group()
