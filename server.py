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
    device = pick_device(args.device)
    cfg = load_config(args.config)
    coord = Coordinator(cfg, device=device, workdir=".")
    if args.local_clients:
        cfg.server.clients = args.local_clients
        coord.logger.log_info("### Application start ###\n")
        specs = coord.local_specs(args.local_clients, args.attacker)
    else:
        coord.listen()
        specs = coord.wait_for_clients(args.register_timeout)
    coord.start(specs)
    coord.run()
    print_with_color("Ok, ready!", "green")
    return 0


if __name__ == "__main__":
    sys.exit(main())
