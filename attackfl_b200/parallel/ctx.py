"""Process-group context: one process per GPU, ``torch.distributed`` (NCCL on GPUs, gloo on CPU).

The reference's "distributed backend" is RabbitMQ + pickle (``server.py:102-108``,
``src/RpcClient.py:174-188``); here NCCL is used for bootstrap and as the debug / baseline collective
path, while the product data path is the fused P2P kernels of ``parallel/symm.py`` + ``csrc/comm.cu``.
Every rank executes the same deterministic control logic (SPMD), so no control messages are needed.
"""
from __future__ import annotations

import os
from dataclasses import dataclass
from typing import Optional

import torch
import torch.distributed as dist


@dataclass
class ParallelContext:
    rank: int = 0
    world: int = 1
    local_rank: int = 0
    device: torch.device = torch.device("cpu")
    group: Optional[object] = None

    @property
    def distributed(self) -> bool:
        return self.world > 1

    # -- client placement: contiguous blocks, rank r owns clients [r*cpg, min(C,(r+1)*cpg))
    def clients_per_rank(self, total: int) -> int:
        return (total + self.world - 1) // self.world

    def local_clients(self, total: int) -> range:
        cpg = self.clients_per_rank(total)
        return range(min(total, self.rank * cpg), min(total, (self.rank + 1) * cpg))

    def owner(self, client: int, total: int) -> int:
        return client // self.clients_per_rank(total)

    # -- collectives (NCCL / gloo); the fused kernels bypass these
    def all_reduce_sum(self, t: torch.Tensor) -> torch.Tensor:
        if self.distributed:
            dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group)
        return t

    def all_reduce_max(self, t: torch.Tensor) -> torch.Tensor:
        if self.distributed:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return t

    def all_gather_rows(self, local: torch.Tensor, total_rows: int) -> torch.Tensor:
        """``local [cpg_or_less, D]`` -> ``[total_rows, D]`` in client order (pads the last rank)."""
        if not self.distributed:
            return local
        cpg = self.clients_per_rank(total_rows)
        pad = torch.zeros(cpg, local.shape[1], dtype=local.dtype, device=local.device)
        pad[:local.shape[0]].copy_(local)
        out = torch.empty(self.world * cpg, local.shape[1], dtype=local.dtype, device=local.device)
        dist.all_gather_into_tensor(out, pad, group=self.group)
        return out[:total_rows]

    def broadcast(self, t: torch.Tensor, src: int = 0) -> torch.Tensor:
        if self.distributed:
            dist.broadcast(t, src=src, group=self.group)
        return t

    def barrier(self) -> None:
        if self.distributed:
            dist.barrier(group=self.group)


def init_from_env(device: Optional[str] = None) -> ParallelContext:
    """Reads RANK / LOCAL_RANK / WORLD_SIZE / MASTER_* (torchrun); falls back to a single process."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    use_cuda = torch.cuda.is_available() and (device is None or str(device).startswith("cuda"))
    dev = torch.device(f"cuda:{local}") if use_cuda else torch.device("cpu")
    if use_cuda:
        torch.cuda.set_device(dev)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29511")
        if use_cuda:
            dist.init_process_group("nccl", rank=rank, world_size=world, device_id=dev)
        else:
            dist.init_process_group("gloo", rank=rank, world_size=world)
    return ParallelContext(rank, world, local, dev, None)


def single(device="cpu") -> ParallelContext:
    return ParallelContext(0, 1, 0, torch.device(device), None)
