"""Python bindings of the persistent training / inference kernels (``afl_train_round``, ``afl_infer``).

Reference parity: replaces ``train_ICU`` (``/root/reference/client.py:75-111``) for all clients of a
rank at once, and the per-batch forward loops of ``src/Validation.py:92-122,178-213``.
"""
from __future__ import annotations

import ctypes
from dataclasses import dataclass
from typing import Optional, Tuple

import torch

from ..models.programs import Program
from . import native
from .native import check, lib_for, ptr, stream_ptr

c_i, c_f, c_i64, c_u64, c_u32 = ctypes.c_int, ctypes.c_float, ctypes.c_int64, ctypes.c_uint64, ctypes.c_uint32


@dataclass
class TrainPlan:
    """How a batch is split: ``G`` blocks per client, each ``M`` tiles of ``S`` samples."""
    G: int
    M: int
    S: int
    chunk: int           # clients per launch (co-residency limit)


def capacity(program: Program, device) -> int:
    if torch.device(device).type != "cuda":
        return 1 << 30
    lib = native.cuda_lib()
    with torch.cuda.device(device):
        cap = int(lib.afl_train_capacity(c_i(program.pool_floats), c_i(program.n_ops),
                                         c_i(program.S if program.chain else 0)))
    if cap <= 0:
        raise RuntimeError("training program does not fit in shared memory: "
                           + lib.afl_last_error().decode(errors="replace"))
    return cap


def make_plan(program: Program, batch: int, n_clients: int, d_pad: int, device, blocks_per_client: int = 0) -> TrainPlan:
    """Choose G (a power of two dividing d_pad/4) so the chunk of clients fills the GPU: as many
    blocks per client as fit (fewer samples per block = shorter critical path) while all of the
    rank's clients stay co-resident in one launch when possible."""
    S = program.S
    tiles = (batch + S - 1) // S
    cap = capacity(program, device)
    g_max = 1
    while g_max * 2 <= tiles and (d_pad // 4) % (g_max * 2) == 0:
        g_max *= 2
    if blocks_per_client > 0:
        G = min(g_max, max(1, blocks_per_client))
        while G & (G - 1):
            G -= 1
    else:
        G = g_max
        while G > 1 and n_clients * G > cap:
            G //= 2
    M = (tiles + G - 1) // G
    chunk = max(1, min(n_clients, cap // G))
    return TrainPlan(G, M, S, chunk)


TILE_CANDIDATES = {"TransformerModel": (8, 16), "RNNModel": (8, 16), "CNNModel": (4,)}


def plan_for(model_name: str, layout, batch: int, n_clients: int, device, *, dropout: bool = True, module=None,
             blocks_per_client: int = 0):
    """Build the training program with the tile size that lets every tile of a batch run on its own block
    (measured on B200: S=8 -> 16 blocks/client is 21-29 % faster than S=16 -> 8 blocks when 128 SMs are free);
    fall back to larger tiles / several tiles per block when the rank hosts many clients."""
    from ..models.programs import build_program
    cands = []
    for S in TILE_CANDIDATES.get(model_name, (16,)):
        prog = build_program(model_name, layout, S=S, train=True, dropout=dropout, module=module).to(device)
        plan = make_plan(prog, batch, n_clients, layout.padded, device, blocks_per_client)
        cands.append((prog, plan))
    # 1) every tile on its own block and all clients co-resident: smallest tile wins (shortest dependent chains);
    # 2) otherwise the blocks are capacity-limited and loop over tiles: fewest, largest tiles win (per-tile
    #    overhead: ~40 barrier intervals), preferring plans that keep all clients in one launch.
    par = [c for c in cands if c[1].M == 1 and c[1].chunk >= n_clients]
    if par:
        return min(par, key=lambda c: c[1].S)
    return min(cands, key=lambda c: (0 if c[1].chunk >= n_clients else 1, c[1].M, -c[1].S))


def train_round(program: Program, plan: TrainPlan, w: torch.Tensor, m: torch.Tensor, v: Optional[torch.Tensor],
                gpart: torch.Tensor, data: torch.Tensor, idx: torch.Tensor, n_data: torch.Tensor,
                client_ids: torch.Tensor, batch: int, epochs: int, *, optimizer: str = "adam", lr: float = 0.004,
                betas=(0.9, 0.999), eps: float = 1e-8, momentum: float = 0.0, clip_proper: bool = False,
                clip: float = 0.0, seed: int = 0, dropout: bool = True,
                scratch: Optional[dict] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """Runs one round of local training for ``w.shape[0]`` clients in place.

    ``w, m, v``: ``[C, d_pad]`` fp32 (``m, v`` must be zeroed by the caller: the reference re-creates
    Adam every round, client.py:78); ``gpart``: ``[>= chunk*G, d_pad]``; ``data``: ``[rows, rec]``;
    ``idx``: int32 ``[C, epochs, max_n]``; ``n_data, client_ids``: int32 ``[C]``.
    Returns ``(loss [C,2] = {sum of sample losses, #NaN}, steps [C])``.
    """
    C, d_pad = w.shape
    dev = w.device
    lib = lib_for(w)
    assert program.ops.device == dev and program.train
    assert idx.dtype == torch.int32 and idx.dim() == 3 and idx.is_contiguous()
    assert gpart.shape[0] >= plan.chunk * plan.G and gpart.shape[1] == d_pad
    loss = torch.zeros(C, 2, dtype=torch.float32, device=dev)
    steps = torch.zeros(C, dtype=torch.int32, device=dev)
    bar = torch.zeros(C, dtype=torch.int32, device=dev)
    gnorm = torch.zeros(C, 2, dtype=torch.float32, device=dev)
    for c0 in range(0, C, plan.chunk):
        c1 = min(C, c0 + plan.chunk)
        rc = lib.afl_train_round(
            ptr(program.ops), c_i(program.n_ops), c_i(program.pool_floats),
            ptr(w[c0:c1]), ptr(m[c0:c1]), ptr(v[c0:c1]) if v is not None else ctypes.c_void_p(0), ptr(gpart),
            c_i64(d_pad), ptr(data), c_i(data.shape[1]), ptr(idx[c0:c1]), c_i64(idx.stride(0)), c_i64(idx.stride(1)),
            ptr(n_data[c0:c1]), ptr(client_ids[c0:c1]), c_i(c1 - c0), c_i(plan.G), c_i(plan.M), c_i(plan.S),
            c_i(batch), c_i(epochs), c_i(0 if optimizer == "adam" else 1), c_f(lr), c_f(betas[0]), c_f(betas[1]),
            c_f(eps), c_f(momentum), c_i(1 if clip_proper else 0), c_f(clip), c_u64(seed & (2 ** 64 - 1)),
            c_i(1 if dropout else 0), ptr(loss[c0:c1]), ptr(steps[c0:c1]), ptr(bar[c0:c1]), ptr(gnorm[c0:c1]),
            c_i(1 if program.chain else 0), stream_ptr(w))
        check(lib, rc, "afl_train_round")
    return loss, steps


def infer(program: Program, w: torch.Tensor, data: torch.Tensor) -> torch.Tensor:
    """Probabilities ``[n_models, n_rows]`` of every row of ``w`` on every record of ``data``."""
    if w.dim() == 1:
        w = w[None]
    w = w.contiguous()
    n_models, d_pad = w.shape
    n_rows = data.shape[0]
    lib = lib_for(w)
    probs = torch.empty(n_models, n_rows, dtype=torch.float32, device=w.device)
    rc = lib.afl_infer(ptr(program.ops), c_i(program.n_ops), c_i(program.pool_floats), ptr(w), c_i64(d_pad),
                       c_i(n_models), ptr(data), c_i(data.shape[1]), c_i64(n_rows), ptr(probs), c_i(program.S),
                       stream_ptr(w))
    check(lib, rc, "afl_infer")
    return probs


def debug_tile(program: Program, w_row: torch.Tensor, data: torch.Tensor, idx: torch.Tensor, n_valid: int,
               inv_b: float, seed: int = 0, client: int = 0, step: int = 0, sample_base: int = 0,
               train: bool = False):
    """CPU twin only: one forward+backward of a single tile; returns (grad row, pool, loss[2])."""
    lib = native.cpu_lib()
    gw = torch.zeros_like(w_row)
    pool = torch.zeros(program.pool_floats, dtype=torch.float32)
    loss = torch.zeros(2, dtype=torch.float32)
    rc = lib.afl_debug_tile(ptr(program.ops), c_i(program.n_ops), c_i(program.pool_floats), ptr(w_row), ptr(gw),
                            ptr(data), c_i(data.shape[1]), ptr(idx), c_i(program.S), c_i(n_valid), c_f(inv_b),
                            c_u64(seed), c_u32(client), c_u32(step), c_i(sample_base), c_i(1 if train else 0),
                            ptr(pool), ptr(loss))
    check(lib, rc, "afl_debug_tile")
    return gw, pool, loss


def dropout_keep(seed: int, client: int, step: int, sample: int, rng: int, elem: int, p: float) -> bool:
    lib = native.cpu_lib()
    return bool(lib.afl_dropout_keep(c_u64(seed), c_u32(client), c_u32(step), c_u32(sample), c_u32(rng),
                                     c_u32(elem), c_f(p)))
