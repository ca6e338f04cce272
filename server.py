#!/usr/bin/env python
"""``python server.py [--device D]`` — coordinator + device runtime (reference: server.py:36-53,838-842).

Reads ``config.yaml`` from the current directory, waits for ``server.clients`` registrations from ``client.py``
processes (or creates them in-process with ``--local-clients N``), then runs the federation: local training of all
client slots, attacks, aggregation / hyper-network update, validation, ``*.pth`` checkpoints, ``app.log``.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from attackfl_b200 import load_config                      # noqa: E402
from attackfl_b200.cli import pick_device, server_parser   # noqa: E402
from attackfl_b200.coordinator import Coordinator          # noqa: E402
from attackfl_b200.log import print_with_color             # noqa: E402


def main(argv=None) -> int:
    args = server_parser().parse_args(argv)
    cfg = load_config(args.config)
    if args.gpus:
        cfg.runtime.gpus = args.gpus
    if args.profile:
        cfg.runtime.profile_round = args.profile
    from attackfl_b200.parallel import launch
    rc = launch.maybe_relaunch(cfg.runtime.gpus)           # one process per GPU (re-exec under torchrun)
    if rc is not None:
        return rc
    ctx = None
    if "RANK" in os.environ:
        from attackfl_b200.parallel import ctx as pctx
        ctx = pctx.init_from_env(args.device)
        device = str(ctx.device)
    else:
        device = pick_device(args.device)
    rank0 = ctx is None or ctx.rank == 0
    coord = Coordinator(cfg, device=device, workdir=".", ctx=ctx)
    specs = None
    if args.local_clients:
        cfg.server.clients = args.local_clients
        coord.logger.log_info("### Application start ###\n")
        specs = coord.local_specs(args.local_clients, args.attacker)
    elif rank0:
        coord.listen()
        specs = coord.wait_for_clients(args.register_timeout)
    if ctx is not None:
        specs = launch.broadcast_specs(ctx, specs)
    fed = coord.start(specs)
    if ctx is not None and ctx.distributed and cfg.runtime.collectives == "fused" and fed.device.type == "cuda":
        from attackfl_b200.parallel import symm
        symm.attach(fed)
    coord.run()
    if rank0:
        print_with_color("Ok, ready!", "green")
    return 0


if __name__ == "__main__":
    sys.exit(main())
