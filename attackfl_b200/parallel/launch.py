"""Rank launcher: one process per GPU (``runtime.gpus``), NCCL bootstrap over 127.0.0.1.

The reference starts one OS process per client plus a broker (``README.md:91-143``); here ``server.py`` re-executes
itself under ``torch.distributed.run`` when ``runtime.gpus > 1`` and the client specs collected by rank 0 are
broadcast to the other ranks, after which every rank runs the same SPMD round loop."""
from __future__ import annotations

import os
import subprocess
import sys
from typing import List, Optional


def maybe_relaunch(gpus: int, argv: Optional[List[str]] = None, port: int = 29561) -> Optional[int]:
    """If ``gpus > 1`` and we are not already under torchrun, re-exec under it and return its exit code."""
    if gpus <= 1 or "RANK" in os.environ:
        return None
    argv = list(sys.argv if argv is None else argv)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={gpus}", "--master-addr",
           "127.0.0.1", "--master-port", str(port)] + argv
    return subprocess.call(cmd)


def broadcast_specs(ctx, specs):
    """Rank 0 collected the registrations; everybody else receives the same list."""
    if not ctx.distributed:
        return specs
    import torch.distributed as dist
    box = [specs if ctx.rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return box[0]
