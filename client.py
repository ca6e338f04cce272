#!/usr/bin/env python
"""``python client.py [--device D] [--attack B --attack_mode M --attack_round R --attack_args a ...]``

Registers one FL client with the coordinator (reference: client.py:19-61,134-142).  The client's model, data
sampling, local training and attack run in its slot of the server's device arena — updates never leave device
memory — so this process only registers, prints what its slot does each round, and exits on STOP.
"""
import os
import sys
import uuid

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

from attackfl_b200 import load_config                                              # noqa: E402
from attackfl_b200.cli import client_parser, pick_device, validate_client_args     # noqa: E402
from attackfl_b200.control import ControlClient                                    # noqa: E402
from attackfl_b200.log import print_with_color                                     # noqa: E402


def main(argv=None) -> int:
    args = client_parser().parse_args(argv)
    validate_client_args(args)
    print(f"Attack: {args.attack}, Mode: {args.attack_mode}")
    cfg = load_config(args.config)
    client_id = uuid.uuid4()
    device = pick_device(args.device)
    print_with_color("[>>>] Client sending registration message to server...", "red")
    print("Client ID:", client_id)
    conn = ControlClient(cfg.runtime.control_host, cfg.runtime.control_port)
    conn.register({"action": "REGISTER", "client_id": str(client_id), "message": "Hello from Client!",
                   "attack": bool(args.attack), "attack_mode": args.attack_mode, "attack_round": args.attack_round,
                   "attack_args": args.attack_args, "device": device})
    for msg in conn.messages():
        action = msg.get("action")
        if action == "START":
            print_with_color(f"[<<<] Client received: {msg.get('message')}", "blue")
        elif action == "UPDATE":
            print_with_color(f"Number of client's data: {msg.get('size')}", "yellow")
            if msg.get("attacking"):
                print_with_color(f"[===] Client is start attacking {args.attack_mode} ...", "red")
            else:
                print_with_color(f"Loss {msg.get('loss')} ", "yellow")
            print_with_color("[>>>] Client sent parameters to server", "red")
        elif action == "STOP":
            print_with_color(f"[<<<] Client received: {msg.get('message')}", "blue")
    conn.close()
    return 0


if __name__ == "__main__":
    sys.exit(main())
