"""Symmetric-memory runtime and the fused multi-GPU round end (K1-fused, ``csrc/comm.cu``).

Replaces the reference's broker traffic (every UPDATE and every START is a pickled ``state_dict`` through
RabbitMQ: ``/root/reference/src/RpcClient.py:111-114``, ``server.py:588-614``) and the NCCL fall-back path
(all-gather of the client rows + replicated aggregation) with one kernel per rank that reads the client rows *in
place* over NVLink and writes the new global model into every rank.

``attach(fed)`` moves the engine's client rows, global model and LIE buffer into
``torch.distributed._symmetric_memory`` tensors (peer-mapped on every rank; PyTorch/NCCL only exchange the
handles), builds the pointer tables and installs ``fed.fused_comm``.
"""
from __future__ import annotations

import ctypes
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..ops import native
from ..ops.native import check, ptr, stream_ptr

MAX_ATT = 4


class FusedComm:
    def __init__(self, fed):
        import torch.distributed._symmetric_memory as symm_mem
        ctx = fed.ctx
        assert ctx.distributed and fed.device.type == "cuda"
        self.fed, self.world, self.rank = fed, ctx.world, ctx.rank
        dev = fed.device
        group = dist.group.WORLD
        self.cpg = ctx.clients_per_rank(fed.C)
        D = fed.D
        self.n_lie = max(1, sum(1 for sp in fed.specs if sp.attack and sp.attack_mode == "LIE"))
        self.rows = symm_mem.empty(self.cpg, D, dtype=torch.float32, device=dev)
        self.glob = symm_mem.empty(D, dtype=torch.float32, device=dev)
        self.lie = symm_mem.empty(self.n_lie, D, dtype=torch.float32, device=dev)
        self.sig = symm_mem.empty(2048, dtype=torch.int32, device=dev)     # [2][32 chunks][16 ranks] flag words
        self.part = symm_mem.empty(1 + 2 * min(self.n_lie, MAX_ATT), D, dtype=torch.float32, device=dev)
        self.rows.zero_(); self.lie.zero_(); self.sig.zero_(); self.part.zero_()
        self.rows[:fed.CL].copy_(fed.rows[:fed.CL])
        self.glob.copy_(fed.global_row)
        self.h_rows = symm_mem.rendezvous(self.rows, group)
        self.h_glob = symm_mem.rendezvous(self.glob, group)
        self.h_lie = symm_mem.rendezvous(self.lie, group)
        self.h_sig = symm_mem.rendezvous(self.sig, group)
        self.h_part = symm_mem.rendezvous(self.part, group)
        torch.cuda.synchronize(dev)
        dist.barrier()
        # the engine now works directly on the symmetric tensors
        fed.rows = self.rows
        fed.global_row = self.glob
        # pointer tables (host arrays of peer-mapped device pointers)
        PtrArr = ctypes.c_void_p * self.world
        self.glob_ptrs = PtrArr(*[int(p) for p in self.h_glob.buffer_ptrs])
        self.lie_ptrs = PtrArr(*[int(p) for p in self.h_lie.buffer_ptrs])
        self.sig_ptrs = PtrArr(*[int(p) for p in self.h_sig.buffer_ptrs])
        self.part_ptrs = PtrArr(*[int(p) for p in self.h_part.buffer_ptrs])
        self.ctrl = torch.zeros(128, dtype=torch.int32, device=dev)            # [2][32] phase counters, then the error word
        self.epoch = 0
        import os
        mc = [int(getattr(h, "multicast_ptr", 0) or 0) for h in (self.h_part, self.h_glob, self.h_lie)]
        # measured on B200: at 2 ranks P2P beats the in-switch reduction (459 vs 308 GB/s at 1 GB: the local copy
        # also crosses NVLink); NVLS is enabled automatically from 4 ranks on, ATTACKFL_NVLS=0/1 overrides
        want = os.environ.get("ATTACKFL_NVLS", "auto")
        self.multicast = all(mc) and (want == "1" or (want == "auto" and self.world >= 4))
        self.mc_part, self.mc_glob, self.mc_lie = mc if self.multicast else (0, 0, 0)
        # hyper mode: row-sharded head, one persistent kernel per rank and round (csrc/hyper_round.cu)
        self.hyper_sharded = False
        nh = getattr(fed, "native_hyper", None)
        if fed.mode == "hyper" and nh is not None and fed.cfg.server.hyper_update == "sequential":
            lib = native.cuda_lib()
            lib.afl_hyper_round_xchg_floats.restype = ctypes.c_int64
            lib.afl_hyper_round_sig_words.restype = ctypes.c_int64
            n_max = fed.C
            self.hx = symm_mem.empty(int(lib.afl_hyper_round_xchg_floats(n_max, self.world, nh.H)), dtype=torch.float32, device=dev)
            self.hs = symm_mem.empty(int(lib.afl_hyper_round_sig_words(n_max, self.world)), dtype=torch.int32, device=dev)
            self.hx.zero_(); self.hs.zero_()
            self.h_hx = symm_mem.rendezvous(self.hx, group)
            self.h_hs = symm_mem.rendezvous(self.hs, group)
            torch.cuda.synchronize(dev)
            dist.barrier()
            sl = D // self.world
            nh.setup_round(self.rank * sl, (self.rank + 1) * sl, self.world, self.rank, n_max,
                           PtrArr(*[int(p) for p in self.h_hx.buffer_ptrs]), PtrArr(*[int(p) for p in self.h_hs.buffer_ptrs]),
                           shard_storage=True)
            self.hyper_sharded = True
            flag = torch.tensor([1.0 if fed.save else 0.0], device=dev)
            dist.all_reduce(flag, op=dist.ReduceOp.MAX)
            self.any_rank_saves = bool(flag.item() > 0)       # checkpoints need the shards gathered: a collective

    def assert_ownership(self, reporting: List[int]) -> None:
        """Debug build of the arena (``ATTACKFL_DEBUG_OWNERSHIP=1``): checks, every round and on every rank,
        the invariants the fused kernels rely on but cannot see — client c lives on rank ``c // cpg`` and nowhere else,
        this rank's view of its own rows is the tensor the trainer writes, the coordinate slices of the ranks partition
        ``[0, D)`` on 128-byte boundaries, a sharded hyper head holds exactly this rank's slice, and every rank enters
        the collective with the same reporting list and epoch (a mismatch would deadlock or mix rounds)."""
        fed = self.fed
        D, W, r = fed.D, self.world, self.rank
        assert fed.local == list(range(r * self.cpg, min(fed.C, (r + 1) * self.cpg))), "clients are not block-distributed"
        assert fed.rows.data_ptr() == self.rows.data_ptr() == int(self.h_rows.buffer_ptrs[r]), "trainer rows are not the symmetric rows"
        assert fed.global_row.data_ptr() == int(self.h_glob.buffer_ptrs[r]), "global row is not the symmetric buffer"
        for c in fed.local:
            assert self.row_ptr(c) == fed.rows[c - fed.local[0]].data_ptr(), f"row of client {c} is not where its owner trains it"
        for c in reporting:
            assert 0 <= c // self.cpg < W, f"client {c} has no owner rank"
        sl = D // W
        assert sl * W == D and sl % 32 == 0, "coordinate slices must partition the row on 128-byte boundaries"
        nh = getattr(fed, "native_hyper", None)
        if self.hyper_sharded and nh is not None:
            assert (nh.lo, nh.hi) == (r * sl, (r + 1) * sl), "hyper head slice is not this rank's coordinate slice"
            if nh.fused.sharded:
                assert nh.head_w.shape[0] == sl and nh.mw.shape[0] == sl and nh.vw.shape[0] == sl
        h = 0
        for c in reporting:
            h = (h * 1000003 + c + 1) % 2147483647
        t = torch.tensor([float(h), -float(h), float(self.epoch), -float(self.epoch)], dtype=torch.float64, device=fed.device)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        assert float(t[0]) == -float(t[1]), "ranks disagree on the reporting clients of this round"
        assert float(t[2]) == -float(t[3]), "ranks disagree on the epoch of the fused collectives"

    def row_ptr(self, client: int) -> int:
        """Device address of a client's row on this rank's view of its owner's arena (peer-mapped)."""
        return int(self.h_rows.buffer_ptrs[client // self.cpg]) + (client % self.cpg) * self.fed.D * 4

    def hyper_gather_full(self, all_ranks: bool = False) -> Optional[Dict[str, torch.Tensor]]:
        """Checkpoint / export only (a collective, off the round path): the row shards of ``W_h`` / ``b_h`` gathered
        into full host tensors on rank 0 (on every rank with ``all_ranks``); the Adam moments stay sharded."""
        nh = self.fed.native_hyper
        out: Dict[str, torch.Tensor] = {}
        for name, t in (("head_w", nh.head_w), ("head_b", nh.head_b)):
            if all_ranks:
                full = torch.empty((self.world * t.shape[0],) + tuple(t.shape[1:]), dtype=t.dtype, device=t.device)
                dist.all_gather_into_tensor(full, t.contiguous())
                out[name] = full.cpu()
            else:
                parts = [torch.empty_like(t) for _ in range(self.world)] if self.rank == 0 else None
                dist.gather(t.contiguous(), parts, dst=0)
                if self.rank == 0:
                    out[name] = torch.cat([p.cpu() for p in parts])
        return out if (all_ranks or self.rank == 0) else None

    def fedavg(self, weights: torch.Tensor, member: Optional[torch.Tensor] = None, z: Optional[torch.Tensor] = None,
               ref: Optional[torch.Tensor] = None, strict_k1: bool = False) -> Optional[torch.Tensor]:
        """Weighted FedAvg over all ranks' rows -> ``fed.global_row`` on every rank and every local client row;
        returns the LIE rows ``[A, D_pad]`` (clone) when ``member`` is given."""
        fed = self.fed
        dev = fed.device
        lib = native.cuda_lib()
        A = 0 if member is None else member.shape[0]
        assert A <= self.n_lie
        lo, hi = fed.local[0] if fed.local else 0, (fed.local[-1] + 1) if fed.local else 0
        w_loc = weights[lo:hi].to(device=dev, dtype=torch.float32).contiguous()
        out = None
        passes = [(0, 0)] if A == 0 else [(a0, min(A, a0 + MAX_ATT)) for a0 in range(0, A, MAX_ATT)]
        for pi, (a0, a1) in enumerate(passes):
            self.epoch += 1
            n = a1 - a0
            if n:
                mem_all = member[a0:a1].to(dtype=torch.float32)
                kc = mem_all.sum(1).to(dev).contiguous()                  # global subset sizes
                mem = mem_all[:, lo:hi].to(dev).contiguous()             # membership of the local clients
                zz = z[a0:a1].to(device=dev, dtype=torch.float32).contiguous()
                lie_ptrs = (ctypes.c_void_p * self.world)(*[int(p) + a0 * fed.D * 4 for p in self.h_lie.buffer_ptrs])
            else:
                mem = kc = zz = None
                lie_ptrs = None
            rc = lib.afl_fedavg_fused(
                ptr(self.rows), ptr(w_loc), ctypes.c_int(fed.CL), ctypes.c_int64(fed.D), ctypes.c_int(self.world),
                ctypes.c_int(self.rank), self.part_ptrs, self.glob_ptrs, lie_ptrs, self.sig_ptrs,
                ctypes.c_uint32(self.epoch), ctypes.c_int(n), ptr(ref), ptr(mem), ptr(kc), ptr(zz),
                ctypes.c_int(1 if strict_k1 else 0),
                ptr(self.rows) if pi == len(passes) - 1 else ctypes.c_void_p(0), ptr(self.ctrl),
                ctypes.c_void_p(self.mc_part), ctypes.c_void_p(self.mc_glob),
                ctypes.c_void_p(self.mc_lie + a0 * fed.D * 4 if self.mc_lie else 0), stream_ptr(self.rows))
            check(lib, rc, "afl_fedavg_fused")
        if A:
            out = self.lie[:A].clone()
        return out

    # ------------------------------------------------------------------ K3 multi-GPU: Gram attacks by K-slice ownership
    def _ensure_gram_buffers(self, n_attackers: int) -> None:
        import torch.distributed._symmetric_memory as symm_mem
        if getattr(self, "craft_sym", None) is not None and self.craft_sym.shape[0] >= n_attackers:
            return
        fed = self.fed
        self.craft_sym = symm_mem.empty(max(n_attackers, 1), fed.D, dtype=torch.float32, device=fed.device)
        self.craft_sym.zero_()
        self.h_craft = symm_mem.rendezvous(self.craft_sym, dist.group.WORLD)
        torch.cuda.synchronize(fed.device)
        dist.barrier()

    def gram_attack(self, slot: int, subset: List[int], mode: str, center: torch.Tensor, paper: bool) -> torch.Tensor:
        """Min-Max / Min-Sum / Opt-Fang for one attacker without gathering any row: every rank handles its K-slice
        (rows pulled over NVLink inside the moments kernel, tcgen05 Gram on the slice), the (k+2)^2 partial Grams are
        summed across ranks, the gamma search runs replicated on the device, and the crafted slice is stored into every
        rank's copy of the attacker's crafted row.  Returns this rank's full crafted row (view of symmetric memory)."""
        from ..ops import flat, gram_native
        fed = self.fed
        dev = fed.device
        lib = native.cuda_lib()
        D, G, k = fed.D, self.world, len(subset)
        sl = D // G
        lo, hi = self.rank * sl, (self.rank + 1) * sl
        ptrs = torch.tensor([int(self.h_rows.buffer_ptrs[c // self.cpg]) + (c % self.cpg) * D * 4 for c in subset],
                            dtype=torch.int64, device=dev)
        ops = torch.empty(k + 2, sl, dtype=torch.float32, device=dev)
        mu, p = torch.empty(sl, dtype=torch.float32, device=dev), torch.empty(sl, dtype=torch.float32, device=dev)
        rc = lib.afl_slice_moments(ptr(ptrs), ctypes.c_int(k), ptr(center), ctypes.c_int64(lo), ctypes.c_int64(hi),
                                   ctypes.c_int64(sl), ctypes.c_int(1 if mode == "Opt-Fang" else 0), ptr(ops), ptr(mu),
                                   ptr(p), stream_ptr(ops))
        check(lib, rc, "afl_slice_moments")
        prec = fed.cfg.attack.gram_precision
        if prec == "fp32" or k + 2 > 256:
            g = ops @ ops.t()                                # configured exact mode (or too many rows for one TMEM tile)
        else:
            g = gram_native.gram_tcgen05(ops, prec)
        dist.all_reduce(g)                                   # (k+2)^2 floats: the only cross-rank traffic besides the rows
        g4 = flat.gamma_search(g, k, "sum" if mode == "Min-Sum" else "max", True, paper)
        dst = (ctypes.c_void_p * G)(*[int(b) + slot * D * 4 for b in self.h_craft.buffer_ptrs])
        rc = lib.afl_craft_scatter(dst, ctypes.c_int(G), ptr(mu), ptr(p), ptr(g4), ctypes.c_int64(lo), ctypes.c_int64(hi),
                                   stream_ptr(ops))
        check(lib, rc, "afl_craft_scatter")
        self._last_gamma = g4
        return self.craft_sym[slot]

    # ------------------------------------------------------------------ small all-reduce through K1-fused (no NCCL)
    def allreduce_small(self, t: torch.Tensor) -> torch.Tensor:
        """Sum of a small fp32 tensor (<= 65,536 elements, e.g. an R x R partial Gram) over all ranks with the fused
        FedAvg kernel itself: every rank contributes one "client row" of weight 1, the result lands in every rank's
        symmetric buffer.  Stays on the stream, no host sync."""
        import torch.distributed._symmetric_memory as symm_mem
        n = t.numel()
        cap = 65536
        assert n <= cap, "allreduce_small: at most 65,536 elements"
        if getattr(self, "g_rows", None) is None:
            dev = self.fed.device
            self.g_rows = symm_mem.empty(1, cap, dtype=torch.float32, device=dev)
            self.g_part = symm_mem.empty(1, cap, dtype=torch.float32, device=dev)
            self.g_glob = symm_mem.empty(cap, dtype=torch.float32, device=dev)
            for b in (self.g_rows, self.g_part, self.g_glob):
                b.zero_()
            hs = [symm_mem.rendezvous(b, dist.group.WORLD) for b in (self.g_part, self.g_glob)]
            symm_mem.rendezvous(self.g_rows, dist.group.WORLD)
            torch.cuda.synchronize(dev)
            dist.barrier()
            PtrArr = ctypes.c_void_p * self.world
            self.g_part_ptrs = PtrArr(*[int(p) for p in hs[0].buffer_ptrs])
            self.g_glob_ptrs = PtrArr(*[int(p) for p in hs[1].buffer_ptrs])
            self.g_one = torch.ones(1, dtype=torch.float32, device=dev)
        lib = native.cuda_lib()
        self.g_rows.view(-1)[:n].copy_(t.reshape(-1))
        d = (n + 4 * self.world * 8 - 1) // (4 * self.world * 8) * (4 * self.world * 8)
        self.epoch += 1
        rc = lib.afl_fedavg_fused(ptr(self.g_rows), ptr(self.g_one), ctypes.c_int(1), ctypes.c_int64(d), ctypes.c_int(self.world),
                                  ctypes.c_int(self.rank), self.g_part_ptrs, self.g_glob_ptrs, None, self.sig_ptrs,
                                  ctypes.c_uint32(self.epoch), ctypes.c_int(0), None, None, None, None, ctypes.c_int(0),
                                  None, ptr(self.ctrl), ctypes.c_void_p(0), ctypes.c_void_p(0), ctypes.c_void_p(0),
                                  stream_ptr(self.g_rows))
        check(lib, rc, "afl_fedavg_fused (small all-reduce)")
        return self.g_glob[:n].view(t.shape).clone()

    def scion_distances(self, reporting: List[int], seed: int, mu_threshold: float = 3.0) -> torch.Tensor:
        """ScionFL scoring on N GPUs (server.py:443-466) without moving a row: every rank quantises its own reporting
        rows to 1-bit codes in a symmetric buffer, the per-client scalars (min, max, bit count) are summed into place by
        the small fused all-reduce, every rank scores its slice of the packed words reading all clients' codes where
        their owner wrote them (peer pointers), and the C + 2 partial sums are all-reduced.  Returns the cosine distances
        ``[len(reporting)]`` (replicated).  Only D / 32 words per client ever cross NVLink."""
        import torch.distributed._symmetric_memory as symm_mem
        from ..ops import flat
        fed = self.fed
        dev = fed.device
        D, words = fed.D, fed.D // 32
        if getattr(self, "sc_bits", None) is None:
            self.sc_bits = symm_mem.empty(self.cpg, words, dtype=torch.int32, device=dev)
            self.sc_bits.zero_()
            self.h_sc = symm_mem.rendezvous(self.sc_bits, dist.group.WORLD)
            torch.cuda.synchronize(dev)
            dist.barrier()
        n = len(reporting)
        pos = {c: j for j, c in enumerate(reporting)}
        mine = [c for c in fed.local if c in pos]                 # ascending ids -> consecutive positions
        scal = torch.zeros(n, 3, dtype=torch.float32, device=dev)
        if mine:
            assert [pos[c] for c in mine] == list(range(pos[mine[0]], pos[mine[0]] + len(mine)))
            loc = [c - fed.local[0] for c in mine]
            rows = fed.rows[loc[0]:loc[-1] + 1] if loc == list(range(loc[0], loc[-1] + 1)) else fed.rows[loc]
            _, counts, minmax = flat.scion_quantize(rows.contiguous(), fed.valid, seed, row0=pos[mine[0]], bits=self.sc_bits)
            scal[pos[mine[0]]:pos[mine[0]] + len(mine), :2] = minmax
            scal[pos[mine[0]]:pos[mine[0]] + len(mine), 2] = counts.to(torch.float32)
        scal = self.allreduce_small(scal)       # also orders every rank's codes before any peer reads them (flag barrier)
        counts, minmax = scal[:, 2].round().to(torch.int32), scal[:, :2]
        d = float(fed.valid.sum().item())
        lo, hi = flat.scion_clipped_levels(counts, minmax, d, mu_threshold)
        slot = {c: k for k, c in enumerate(mine)}
        ptrs = []
        for c in reporting:                                        # slot of a client inside its owner's code buffer
            owner = c // self.cpg
            first = next(j for j, cc in enumerate(reporting) if cc // self.cpg == owner)
            ptrs.append(int(self.h_sc.buffer_ptrs[owner]) + (pos[c] - first) * words * 4)
        wl = words // self.world
        part = flat.scion_scores_ptrs(ptrs, words, self.rank * wl, (self.rank + 1) * wl if self.rank < self.world - 1 else words,
                                      fed.valid, torch.stack([lo, hi], 1).to(dev))
        sums = self.allreduce_small(part)
        return flat.scion_distances_from_sums(sums, counts, lo, hi, d)

    def check_error(self, raise_on_error: bool = True) -> int:
        """Error word of the fused kernels (non-zero: a cross-GPU flag barrier timed out, the outputs are stale).  The
        word is summed over ranks so that every rank takes the same decision."""
        err = self.ctrl[64:65].to(torch.float32)
        dist.all_reduce(err)
        code = int(err.item())
        if code and raise_on_error:
            raise RuntimeError("fused collective timed out waiting for a peer rank")
        return code

    def reset_error(self) -> None:
        """Re-arms the kernels after a reported time-out (clears the local grid counter, go flag and error word)."""
        self.ctrl.zero_()


def attach(fed) -> FusedComm:
    fc = FusedComm(fed)
    fed.fused_comm = fc
    return fc
