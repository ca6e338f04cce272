"""Command-line surface of ``server.py`` / ``client.py`` (same flags as the reference).

Reference parity: ``/root/reference/server.py:36-53`` (``--device``) and ``client.py:19-38`` (``--device``,
``--attack``, ``--attack_mode {Random,Min-Max,Min-Sum,Opt-Fang,LIE}``, ``--attack_round``, ``--attack_args``).  The
reference declares ``--attack`` with ``type=bool``, so *any* non-empty string — including ``False`` — enables the
attack; here ``True/False/1/0/yes/no`` parse properly and other non-empty strings stay truthy.
"""
from __future__ import annotations

import argparse
import sys
from typing import Optional

from .config import ATTACK_MODES


def parse_bool(v) -> bool:
    if isinstance(v, bool):
        return v
    if v is None:
        return False
    s = str(v).strip().lower()
    if s in ("true", "1", "yes", "y", "on"):
        return True
    if s in ("false", "0", "no", "n", "off", ""):
        return False
    return True                      # reference quirk kept for unknown non-empty strings


def pick_device(arg: Optional[str]) -> str:
    """``--device`` or ``cuda`` if available else ``cpu`` (server.py:42-53, client.py:50-61)."""
    import torch
    if arg is None:
        if torch.cuda.is_available():
            print(f"Using device: {torch.cuda.get_device_name('cuda')}")
            return "cuda"
        print("Using device: CPU")
        return "cpu"
    print(f"Using device: {arg}")
    return arg


def server_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Federated learning framework with controller.")
    p.add_argument("--device", type=str, required=False, help="Device of server")
    # extensions (not in the reference)
    p.add_argument("--config", type=str, default="config.yaml", help="config file (default: ./config.yaml)")
    p.add_argument("--local-clients", type=int, default=0,
                   help="register this many simulated clients in-process instead of waiting for client.py")
    p.add_argument("--attacker", action="append", default=[], metavar="IDX:MODE:ROUND[:ARGS]",
                   help="with --local-clients: mark client IDX as an attacker, e.g. 7:LIE:2:0.74")
    p.add_argument("--register-timeout", type=float, default=None, help="seconds to wait for registrations")
    p.add_argument("--gpus", type=int, default=0, help="ranks to launch (one per GPU); default runtime.gpus")
    p.add_argument("--profile", type=int, default=0, metavar="ROUND",
                   help="bracket round attempt ROUND with cudaProfilerStart/Stop (ncu/nsys capture-range hook)")
    return p


def client_parser() -> argparse.ArgumentParser:
    p = argparse.ArgumentParser(description="Split learning framework")
    p.add_argument("--device", type=str, required=False, help="Device of client")
    p.add_argument("--attack", type=parse_bool, required=False, default=False,
                   help="Set to True to enable attack mode, False otherwise.")
    p.add_argument("--attack_mode", type=str, choices=list(ATTACK_MODES),
                   help="Mode of operation when the attack is enabled (e.g., Random or Min-Max, Min-Sum ...).")
    p.add_argument("--attack_round", type=int, help="Client attack at round.")
    p.add_argument("--attack_args", type=float, nargs="+", required=False,
                   help="A list of attack args. Use space to separate values (e.g., 0.1 0.2 0.5).")
    p.add_argument("--config", type=str, default="config.yaml", help="config file (default: ./config.yaml)")
    return p


def validate_client_args(args) -> None:
    """client.py:32-38 (``attack_round 0`` counts as missing, like ``not 0`` in the reference)."""
    if args.attack and not args.attack_mode:
        print("Error: --attack_mode is required when --attack is True.")
        sys.exit()
    if args.attack and not args.attack_round:
        print("Error: --attack_round is required when --attack is True.")
        sys.exit()
