"""Federation engine: the device runtime + round logic of one rank (public Python API).

Replaces, in one SPMD program per GPU: the reference's server round loop (``/root/reference/server.py``:
``on_request`` :205-275, ``process_consumer`` :277-567, ``notify_clients`` :569-624, ``train_hyper``
:637-678, ``train_FLTrust`` :703-743, ``avg_all_parameters`` :751-775), the client state machine
(``src/RpcClient.py``: ``response_message`` :64-117, ``malicious_training`` :119-145,
``genuine_training`` :147-172) and local training (``client.py:66-131``).

Design: client models are rows of a ``[C_local, D_pad]`` fp32 arena; "sending an UPDATE" is a
no-op, "receiving START" is the broadcast epilogue of the aggregation kernel (fedavg) or the hyper-network
emit kernel (hyper).  All ranks run the same deterministic control logic, so the only cross-rank traffic is
tensor data, moved by the fused kernels (``parallel/symm.py``) or by NCCL (debug / baseline path).
Only per-client ``size`` / ``result`` scalars and the round metric ever reach the host.
"""
from __future__ import annotations

import os
import random
import time
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence

import numpy as np
import torch

from . import aggregators as agg
from . import attacks as atk
from . import models
from .config import Config
from .data import as_har, load_test, load_train, make_round_indices
from .detection import EmbeddingMonitor
from .layout import FlatLayout
from .log import MetricsWriter, NullLogger, print_with_color
from .models.hyper import FusedHyperNet, HyperNetwork
from .models.programs import SUPPORTED as NATIVE_MODELS, build_program
from .parallel.ctx import ParallelContext, single
from .utils.timers import PhaseTimer, Watchdog
from .validation import Validator

GRAM_MODES = ("Min-Max", "Min-Sum", "Opt-Fang")


@dataclass
class ClientSpec:
    """What ``client.py`` registers: identity + attack flags (client.py:19-38,134-142)."""
    index: int
    attack: bool = False
    attack_mode: Optional[str] = None
    attack_round: int = 0
    attack_args: Optional[List[float]] = None
    device: Optional[str] = None
    client_id: Optional[str] = None


@dataclass
class RoundResult:
    round: int
    ok: bool
    metric: Optional[float]
    losses: List[float]
    sizes: List[int]
    attackers: List[int]
    timings_ms: Dict[str, float]
    info: Dict[str, object] = field(default_factory=dict)


class Federation:
    def __init__(self, cfg: Config, specs: Sequence[ClientSpec], device: Optional[str] = None,
                 ctx: Optional[ParallelContext] = None, train_set=None, test_set=None,
                 logger=None, workdir: str = ".", save: bool = True, verbose: bool = True):
        self.cfg = cfg
        self.specs = list(specs)
        self.C = len(self.specs)
        if ctx is None:
            if device is None:
                device = "cuda" if torch.cuda.is_available() else "cpu"
            ctx = single(device)
        self.ctx = ctx
        self.device = ctx.device
        self.workdir, self.save, self.verbose = workdir, save, verbose
        self.logger = logger or NullLogger()
        s, l = cfg.server, cfg.learning
        self.mode, self.model_name, self.data_name = s.mode, s.model, s.data_name
        self.py_rng = random.Random(s.random_seed)            # server.py:96-97
        self.client_rng = [random.Random((s.random_seed or 0) * 7919 + 13 * c + 1) for c in range(self.C)]
        self.torch_gen = torch.Generator(device="cpu").manual_seed(int(s.random_seed or 0) + 1234)
        # per-rank stream for the clients' private data sampling (clients on different ranks must not draw
        # the same subsets); everything that decides control flow uses the rank-independent generators above
        self.dev_gen = torch.Generator(device=self.device).manual_seed(int(s.random_seed or 0) + 4321 + 7919 * ctx.rank)

        # ---- model, layout, programs
        torch.manual_seed(int(s.random_seed or 0))
        self.module = models.build(self.model_name)
        self.layout = FlatLayout.from_module(self.module)
        self.D = self.layout.padded
        self.valid = self.layout.valid_mask(self.device).to(torch.float32)
        backend = cfg.runtime.backend
        native_ok = self.model_name in NATIVE_MODELS and self.data_name == "ICU"
        if backend == "auto":
            backend = "native" if native_ok else "torch"
        if backend == "native" and not native_ok:
            raise ValueError(f"no native training program for {self.model_name}/{self.data_name}")
        self.backend = backend
        self.train_prog = self.infer_prog = None
        # HAR: native trainer (tcgen05 linears + attention / LayerNorm kernels) on CUDA unless backend: torch is forced
        self.har_trainer = None
        if (self.data_name == "HAR" and self.model_name == "TransformerClassifier" and self.device.type == "cuda"
                and cfg.runtime.backend in ("auto", "native")):
            from .ops.har import HarNativeTrainer
            if cfg.learning.batch_size <= 512:                   # activation buffers: ~4 MB per sample of the batch
                self.har_trainer = HarNativeTrainer(self.layout, self.device, max_batch=max(cfg.learning.batch_size, 1))
        if backend == "native":
            from .ops import train as T
            self.T = T
            for s_inf in (32, 16, 8, 4):                  # largest inference tile whose pool fits in shared memory
                prog = build_program(self.model_name, self.layout, S=s_inf, train=False, module=self.module)
                if prog.smem_bytes() <= 200 * 1024:
                    break
            self.infer_prog = prog.to(self.device)

        # ---- data (device resident)
        if train_set is None:
            train_set, self.train_src = load_train(self.data_name, cfg.runtime.synthetic_train_rows, s.random_seed or 0)
        else:
            self.train_src = "provided"
        if test_set is None:
            test_set, self.test_src = load_test(self.data_name, cfg.runtime.synthetic_test_rows, s.random_seed or 0)
        else:
            self.test_src = "provided"
        self._set_data(train_set, test_set)
        if s.num_data_range[1] > self.n_train:
            raise ValueError(f"num-data-range upper bound {s.num_data_range[1]} exceeds the training set "
                             f"({self.n_train} rows); random.sample would fail in the reference as well")

        # ---- arena
        self.local = list(ctx.local_clients(self.C))
        self.CL = len(self.local)
        z = lambda *shape: torch.zeros(*shape, dtype=torch.float32, device=self.device)
        self.rows = z(max(self.CL, 1), self.D)
        self.m = z(max(self.CL, 1), self.D)
        self.v = z(max(self.CL, 1), self.D)
        self.global_row = z(self.D)
        self.have_global = False
        self.crafted: Dict[int, torch.Tensor] = {}           # attacker -> crafted row for its next round
        self.crafted_own: Dict[int, bool] = {}               # attacker returns its received weights (k <= 1)
        self.plan = None
        self.gpart = None
        if backend == "native":
            self.train_prog, self.plan = self.T.plan_for(self.model_name, self.layout, l.batch_size, max(self.CL, 1),
                                                         self.device, dropout=l.dropout, module=self.module,
                                                         blocks_per_client=cfg.runtime.cta_per_client)
            self.gpart = z(self.plan.chunk * self.plan.G, self.D)
        # Tensor-core trainer (csrc/train_tc.cu): one CTA per client, so it wins as soon as the interpreter cannot give
        # every client 16 blocks (measured on B200: 88 us per optimiser step for any client count up to the SM count,
        # against 87 us for the interpreter at 8 clients and 943 us at 128).
        self.trainer = "ffma"
        self.tc_stash = None
        want = os.environ.get("ATTACKFL_TRAINER", cfg.runtime.trainer)
        if backend == "native" and self.device.type == "cuda" and want != "ffma":
            from .ops import train_tc as TC
            ok = TC.supported(self.model_name, self.layout, l.batch_size, self.module)
            if want == "tcgen05" and not ok:
                raise ValueError("runtime.trainer: tcgen05 supports the default-width TransformerModel with batch <= 128")
            if ok and (want == "tcgen05" or self.plan.G < 16):
                self.trainer = "tcgen05"
                self.TC = TC
                self.tc_cache: Dict = {}
                self.tc_gw = z(max(self.CL, 1), self.D)
                self.tc_stash = z(max(self.CL, 1), TC.stash_floats_per_client(self.device))

        # ---- coordinator state (server.py:105-134)
        self.num_round = s.num_round
        self.remaining = s.num_round
        self.attempt = 0
        self.failures_in_a_row = 0
        self.selected = list(range(self.C))
        self.training_round = [0] * self.C                   # RpcClient.training_round
        self.sizes = [0] * self.C
        self.genuine_prev: Optional[List[int]] = None        # indices of last round's genuine reporters
        self.timer = PhaseTimer(self.device, nvtx=cfg.runtime.nvtx)
        self.watchdog = Watchdog(cfg.runtime.round_timeout_s)
        self.validator = Validator(self.data_name, self.logger if ctx.rank == 0 else None, quiet=ctx.rank != 0,
                                   native=self.backend == "native")
        self.metrics = MetricsWriter(os.path.join(workdir, cfg.runtime.metrics_file) if (save and ctx.rank == 0) else None)
        self.history: List[RoundResult] = []
        self.fused_comm = None                               # set by parallel.symm.attach()

        # ---- initial parameters
        init_row = self.layout.flatten(self.module.state_dict(), device=self.device)
        self.init_row = init_row
        pth = os.path.join(workdir, f"{self.model_name}.pth")
        if self.mode == "hyper":
            torch.manual_seed(int(s.random_seed or 0))
            target = models.build(self.model_name)
            hpth = os.path.join(workdir, f"{self.model_name}_hyper_{self.C}.pth")
            if s.load_parameters and not os.path.exists(hpth) and os.path.exists(pth):
                print_with_color(f"Load state dict from original model: {pth}", "yellow")
                target.load_state_dict(torch.load(pth, map_location="cpu", weights_only=True))
            hnet = HyperNetwork(target, self.C, 8, 100, False, 2)          # server.py:799-801
            self.hyper = FusedHyperNet.from_module(hnet, self.layout, device=self.device)
            if s.load_parameters and os.path.exists(hpth):                  # working resume
                print_with_color(f"Load state dict from hyper model: {hpth}", "yellow")
                self.hyper.import_state_dict(torch.load(hpth, map_location="cpu", weights_only=True))
            self.native_hyper = None
            if backend == "native":
                from .ops.flat import NativeHyper
                self.native_hyper = NativeHyper(self.hyper)
            self.monitor = None
            hd = s.hyper_detection
            if hd.enable:
                self.monitor = EmbeddingMonitor(self.C, hd.cosine_search if hd.honour_cosine_search else 5,
                                                hd.start_round, hd.n_components, hd.eps, hd.min_samples,
                                                os.path.join(workdir, "all_embeddings.npy") if save else "")
        else:
            self.hyper = None
            if s.load_parameters and os.path.exists(pth):
                self.global_row.copy_(self.layout.load_pth(pth, device=self.device))
                self.have_global = True
            elif s.common_init:
                self.global_row.copy_(init_row)
                self.have_global = True
            else:
                # reference behaviour (Q4): every client starts from its own random init
                for j, c in enumerate(self.local):
                    torch.manual_seed(int(s.random_seed or 0) * 1000 + c + 1)
                    self.rows[j].copy_(self.layout.flatten(models.build(self.model_name).state_dict(), device=self.device))
        if self.mode == "FLTrust":
            self.server_row = z(self.D)

    # ================================================================================== data
    def _set_data(self, train_set, test_set) -> None:
        if self.data_name == "ICU":
            self.train_dev = train_set.packed().to(self.device)
            self.test_dev = test_set.packed().to(self.device)
            self.test_labels = self.test_dev[:, -1].contiguous()
        else:
            tr, te = as_har(train_set), as_har(test_set)
            self.train_x, self.train_y = tr.x.to(self.device), tr.y.to(self.device)
            self.test_x, self.test_y = te.x.to(self.device), te.y.to(self.device)
            self.train_dev = self.train_x
        self.n_train = self.train_dev.shape[0]
        self.host_train = None

    def set_host_train(self, pinned: torch.Tensor) -> None:
        """End-to-end mode: the training pool lives in pinned host memory and is copied to the device at the
        start of every round (H2D bytes = ``pinned.numel()*4``), like a host-resident DataLoader would."""
        assert pinned.shape == self.train_dev.shape
        self.host_train = pinned

    # ================================================================================== helpers
    def _log(self, text: str, color: str = "yellow") -> None:
        if self.verbose and self.ctx.rank == 0:
            print_with_color(text, color)

    def _row_of(self, c: int) -> torch.Tensor:
        return self.rows[c - self.local[0]]

    def _all_rows(self) -> torch.Tensor:
        """All clients' rows in client order (identity on one rank; all-gather otherwise)."""
        return self.ctx.all_gather_rows(self.rows[:self.CL], self.C)

    def _gram_fn(self):
        if self.device.type == "cuda":
            from .ops import gram
            prec = self.cfg.attack.gram_precision
            return lambda a: gram.gram(a, precision=prec)
        return None

    # ================================================================================== START
    def _distribute(self, starting: List[int]) -> None:
        if self.mode == "hyper":
            with self.timer.phase("hyper_emit"):
                loc = [c for c in starting if c in self.local]
                if self._hyper_sharded():
                    # every rank emits its row shard of every starting client's model straight into the client's row
                    if not self._rows_hold_theta:
                        self.native_hyper.emit_rows(starting, [self.fused_comm.row_ptr(c) for c in starting])
                        torch.cuda.synchronize(self.device)
                        self.ctx.barrier()
                elif loc:
                    if self.native_hyper is not None:
                        theta = self.native_hyper.emit(loc)
                    else:
                        theta = self.hyper.emit_torch(torch.tensor(loc, device=self.device))
                    for j, c in enumerate(loc):
                        self._row_of(c).copy_(theta[j])
        elif self.have_global and not self._rows_hold_global:
            self.rows[:self.CL].copy_(self.global_row[None].expand(self.CL, self.D))
        self._rows_hold_global = False
        self._rows_hold_theta = False

    _rows_hold_global = False
    _rows_hold_theta = False

    def _hyper_sharded(self) -> bool:
        return self.fused_comm is not None and getattr(self.fused_comm, "hyper_sharded", False)

    # ================================================================================== local training
    def _cpu_slots(self) -> List[int]:
        """Local slots whose client registered with ``--device cpu`` while this rank trains on a GPU
        (``/root/reference/client.py:50-61``: every client picks its own training device)."""
        if self.device.type != "cuda":
            return []
        return [j for j, c in enumerate(self.local) if str(self.specs[c].device or "").startswith("cpu")]

    def _round_indices(self, sizes_local: List[int]) -> torch.Tensor:
        """int32 ``[CL, epochs, max_size]``: the records every local client visits this round, per epoch
        (``random.sample`` + ``DataLoader(shuffle=True)``, ``/root/reference/src/RpcClient.py:166,169``).  Default: keyed
        permutations in one native launch (``ops.flat.round_indices``; the same indices on every device and for every
        trainer); ``runtime.index-rng: torch`` keeps the argsort-of-random-keys generator on the run's device generator."""
        l = self.cfg.learning
        sizes = [max(s, 1) for s in sizes_local]
        if self.cfg.runtime.index_rng == "torch":
            return make_round_indices(self.n_train, sizes, l.epoch, self.dev_gen, self.device)
        from .ops import flat
        st = torch.tensor(sizes, dtype=torch.int32, device=self.device)
        cids = torch.tensor(self.local, dtype=torch.int32, device=self.device)
        seed = (int(self.cfg.server.random_seed or 0) * 0x9E3779B97F4A7C15 + self.attempt * 0xD1B54A32D192ED03 + 0x1234) & (2 ** 64 - 1)
        return flat.round_indices(self.n_train, st, cids, l.epoch, max(sizes), seed)

    def _train_native(self, sizes_local: List[int]) -> torch.Tensor:
        l = self.cfg.learning
        T = self.T
        cpu_slots = self._cpu_slots()
        sizes_dev = [0 if j in cpu_slots else n for j, n in enumerate(sizes_local)]
        n_data = torch.tensor(sizes_dev, dtype=torch.int32, device=self.device)
        idx = self._round_indices(sizes_local)
        self.m[:self.CL].zero_()
        self.v[:self.CL].zero_()
        cids = torch.tensor(self.local, dtype=torch.int32, device=self.device)
        clip_proper = l.clip_mode == "proper" and l.clip_grad_norm > 0
        seed = (int(self.cfg.server.random_seed or 0) << 20) ^ (self.attempt * 2654435761 & 0xFFFFFFFF)
        kw = dict(optimizer=l.optimizer, lr=l.learning_rate, momentum=l.momentum, clip_proper=clip_proper,
                  clip=l.clip_grad_norm, seed=seed, dropout=l.dropout)
        if self.trainer == "tcgen05":
            loss, steps = self.TC.train_round_tc(self.layout, self.rows[:self.CL], self.m[:self.CL], self.v[:self.CL],
                                                 self.tc_gw, self.tc_stash, self.train_dev, idx, n_data, cids,
                                                 l.batch_size, l.epoch, cache=self.tc_cache, **kw)
        else:
            loss, steps = T.train_round(self.train_prog, self.plan, self.rows[:self.CL], self.m[:self.CL],
                                        self.v[:self.CL], self.gpart, self.train_dev, idx, n_data, cids, l.batch_size,
                                        l.epoch, **kw)
        if cpu_slots:
            # clients that asked for the CPU train on the CPU twin of the same kernel sources (same Philox streams, same
            # sample indices); their rows are copied out and back in
            if getattr(self, "_cpu_prog", None) is None:
                from .models.programs import build_program
                self._cpu_prog = build_program(self.model_name, self.layout, train=True, dropout=l.dropout, module=self.module)
                self._cpu_data = self.train_dev.cpu()
            sel = torch.tensor(cpu_slots, dtype=torch.long, device=self.device)
            w = self.rows[sel].cpu()
            m, v = torch.zeros_like(w), torch.zeros_like(w)
            plan = T.TrainPlan(1, (l.batch_size + self._cpu_prog.S - 1) // self._cpu_prog.S, self._cpu_prog.S, len(cpu_slots))
            gp = torch.zeros(len(cpu_slots), self.D)
            closs, _ = T.train_round(self._cpu_prog, plan, w, m, v, gp, self._cpu_data, idx[sel].cpu().contiguous(),
                                     torch.tensor([sizes_local[j] for j in cpu_slots], dtype=torch.int32),
                                     torch.tensor([self.local[j] for j in cpu_slots], dtype=torch.int32), l.batch_size,
                                     l.epoch, **kw)
            self.rows[sel] = w.to(self.device)
            loss[sel] = closs.to(self.device)
        return loss

    def _train_torch(self, sizes_local: List[int]) -> torch.Tensor:
        """Fallback / oracle path: the reference's loop per client with torch modules on arena views and
        the fused flat optimiser kernel (also the HAR path, client.py:114-131)."""
        from .ops import flat
        l = self.cfg.learning
        loss_out = torch.zeros(self.CL, 2, device=self.device)
        idx = self._round_indices(sizes_local)
        mod = self.module.to(self.device)
        for j in range(self.CL):
            n = sizes_local[j]
            if n == 0:
                continue
            row = self.rows[j]
            params = {k: v.detach().clone().requires_grad_(True) for k, v in self.layout.views(row).items()}
            names = [k for k, _ in mod.named_parameters()]
            self.m[j].zero_(); self.v[j].zero_()
            mod.train(l.dropout)
            step = 0
            for e in range(l.epoch):
                for s0 in range(0, n, l.batch_size):
                    ii = idx[j, e, s0:min(n, s0 + l.batch_size)].long()
                    if self.data_name == "ICU":
                        if ii.numel() == 1:
                            continue
                        rec = self.train_dev[ii]
                        out = torch.func.functional_call(mod, params, (rec[:, :7], rec[:, 7:23]))
                        loss = torch.nn.functional.binary_cross_entropy(out, rec[:, 23:24])
                    else:
                        out = torch.func.functional_call(mod, params, (self.train_x[ii],))
                        loss = torch.nn.functional.cross_entropy(out, self.train_y[ii])
                    if torch.isnan(loss):
                        loss_out[j, 1] += 1
                        break
                    grads = torch.autograd.grad(loss, [params[k] for k in names], allow_unused=True)
                    gd = {k: torch.zeros_like(v) for k, v in params.items() if k not in names}   # buffers (HAR pe.pe)
                    gd.update({k: (gi if gi is not None else torch.zeros_like(params[k])) for k, gi in zip(names, grads)})
                    g = self.layout.flatten(gd, device=self.device)
                    gs = None
                    if l.clip_mode == "proper" and l.clip_grad_norm > 0:
                        gs = torch.clamp(l.clip_grad_norm / (g.norm() + 1e-6), max=1.0).reshape(1)
                    step += 1
                    if l.optimizer == "adam":
                        flat.adam_flat(row, g, self.m[j], self.v[j], l.learning_rate, step, gscale=gs)
                    else:
                        flat.sgdm_flat(row, g, self.m[j], l.learning_rate, l.momentum, step == 1, gscale=gs)
                    with torch.no_grad():
                        for k, v in self.layout.views(row).items():
                            params[k].copy_(v)
                    loss_out[j, 0] += loss.detach() * ii.numel()
        return loss_out

    def _train_har_native(self, sizes_local: List[int]) -> torch.Tensor:
        """HAR local training (``train_HAR``, client.py:114-131) with the native step: per batch one forward + backward
        through ``ops.har`` (gradients written in arena layout) and one fused flat Adam / SGD launch; no host sync."""
        from .ops import flat
        l = self.cfg.learning
        tr = self.har_trainer
        loss_out = torch.zeros(self.CL, 2, device=self.device)
        idx = self._round_indices(sizes_local)
        seed = (int(self.cfg.server.random_seed or 0) << 20) ^ (self.attempt * 2654435761 & 0xFFFFFFFF)
        if not hasattr(self, "_har_grad"):
            self._har_grad = torch.zeros(self.D, device=self.device)
        g = self._har_grad
        x_all = self.train_x.reshape(self.train_x.shape[0], -1)
        for j in range(self.CL):
            n = sizes_local[j]
            if n == 0:
                continue
            row = self.rows[j]
            self.m[j].zero_(); self.v[j].zero_()
            tr.out3.zero_()
            step = 0
            for e in range(l.epoch):
                for s0 in range(0, n, l.batch_size):
                    ii = idx[j, e, s0:min(n, s0 + l.batch_size)].long()
                    step += 1
                    tr.grad_step(row, g, x_all[ii], self.train_y[ii], seed=seed, client=self.local[j], step=step,
                                 dropout=l.dropout)
                    gs = None
                    if l.clip_mode == "proper" and l.clip_grad_norm > 0:
                        gs = torch.clamp(l.clip_grad_norm / (g.norm() + 1e-6), max=1.0).reshape(1)
                    if l.optimizer == "adam":
                        flat.adam_flat(row, g, self.m[j], self.v[j], l.learning_rate, step, gscale=gs)
                    else:
                        flat.sgdm_flat(row, g, self.m[j], l.learning_rate, l.momentum, step == 1, gscale=gs)
            loss_out[j, 0] = tr.out3[0]
            loss_out[j, 1] = tr.out3[1]
        return loss_out

    # ================================================================================== attacks
    def _prepare_crafts(self, rows_all: Optional[torch.Tensor], genuine: List[int], center: Optional[torch.Tensor],
                        defer_lie: bool = False) -> Dict:
        """End of round r: every attacker samples its subset of this round's genuine updates
        (``random.sample(..., max(int(rate*n),1))``, server.py:596-600) and its round-(r+1) row is crafted
        now, from the rows still in the arena (SURVEY §7.1 decision 2)."""
        info: Dict[str, object] = {}
        attackers = [sp for sp in self.specs if sp.attack and sp.index in self.selected]
        if not attackers or not genuine:
            return info
        k = max(int(self.cfg.server.genuine_rate * len(genuine)), 1)
        subsets = {sp.index: self.py_rng.sample(genuine, k) for sp in attackers}
        a_cfg = self.cfg.attack
        native = self.backend == "native"
        lie = [sp for sp in attackers if sp.attack_mode == "LIE"]
        if lie and defer_lie:
            # the moments are computed inside the K1 pass of the aggregation (one GPU: the same launch as FedAvg;
            # N GPUs: the K1-fused pass over peer memory)
            member = torch.zeros(len(lie), self.C)
            for q, sp in enumerate(lie):
                member[q, subsets[sp.index]] = 1.0
            info["_lie_specs"] = lie
            info["_lie_member"] = member
            info["_lie_z"] = torch.tensor([float((sp.attack_args or [0.74])[0]) for sp in lie])
            if rows_all is not None:
                info["_lie_ref"] = (center if center is not None else rows_all[genuine[0]]).contiguous()
        elif lie:
            if native:
                from .ops import flat
                member = torch.zeros(len(lie), rows_all.shape[0])
                for q, sp in enumerate(lie):
                    member[q, subsets[sp.index]] = 1.0
                z = torch.tensor([float((sp.attack_args or [0.74])[0]) for sp in lie])
                ref = center if center is not None else rows_all[genuine[0]]
                _, _, _, rows = flat.fedavg(rows_all, torch.zeros(rows_all.shape[0], device=self.device), out=None,
                                            member=member, z=z, ref=ref.contiguous(), want_lie=True,
                                            strict_k1=a_cfg.strict_reference)
                for q, sp in enumerate(lie):
                    self.crafted[sp.index] = rows[q]
                    self.crafted_own[sp.index] = False
            else:
                for sp in lie:
                    self.crafted[sp.index] = atk.lie_attack(rows_all[subsets[sp.index]], float((sp.attack_args or [0.74])[0]),
                                                            a_cfg.strict_reference)
                    self.crafted_own[sp.index] = False
        gram_specs = [sp for sp in attackers if sp.attack_mode in GRAM_MODES]
        for sp in gram_specs:
            if len(subsets[sp.index]) <= 1:
                self.crafted_own[sp.index] = True                # "return state_dict" (Utils.py:102-103)
        live = [sp for sp in gram_specs if len(subsets[sp.index]) > 1]
        batched = (live and native and self.device.type == "cuda" and a_cfg.metric == "flat"
                   and len(genuine) + len(live) <= 256 and (rows_all is not None or self.fused_comm is not None))
        if batched:
            # ONE Gram for all attackers (csrc/gram_multi.cu): one pass over the genuine rows, one tcgen05 Gram of
            # [genuine rows ; one direction row per attacker], all bisections in one launch, all crafted rows in one launch
            from .ops import flat
            pos = {c: j for j, c in enumerate(genuine)}
            subs = [[pos[c] for c in subsets[sp.index]] for sp in live]
            modes = [sp.attack_mode for sp in live]
            if rows_all is not None:
                cen = center if center is not None else rows_all[genuine].mean(0)
                out, gam = flat.multi_gram_attack([rows_all[c].data_ptr() for c in genuine], cen.contiguous(), subs, modes,
                                                  self.D, paper=a_cfg.search == "paper", precision=a_cfg.gram_precision)
                for a, sp in enumerate(live):
                    self.crafted[sp.index] = out[a]
            else:
                # multi-GPU: K-slice ownership, genuine rows read in place over NVLink, partial Grams summed by K1-fused,
                # crafted slices stored into every rank's copy of the crafted rows
                fc = self.fused_comm
                fc._ensure_gram_buffers(len(live))
                sl = self.D // fc.world
                cen = center if center is not None else self.global_row
                dst = [[int(b) + a * self.D * 4 for b in fc.h_craft.buffer_ptrs] for a in range(len(live))]
                _, gam = flat.multi_gram_attack([fc.row_ptr(c) for c in genuine], cen.contiguous(), subs, modes, self.D,
                                                lo=fc.rank * sl, hi=(fc.rank + 1) * sl, paper=a_cfg.search == "paper",
                                                precision=a_cfg.gram_precision if len(genuine) + len(live) <= 256 else "fp32",
                                                dst_ptrs=dst, reduce_gram=fc.allreduce_small)
                torch.cuda.synchronize(self.device)
                self.ctx.barrier()                                # every rank's slice has landed in every crafted row
                for a, sp in enumerate(live):
                    self.crafted[sp.index] = fc.craft_sym[a].clone()
            for a, sp in enumerate(live):
                info[f"gamma_{sp.index}"] = gam[a]
                self.crafted_own[sp.index] = False
            return info
        if gram_specs and rows_all is None:
            # fused multi-GPU path, one attacker at a time (more than 256 Gram rows): K-slice ownership (parallel/symm.py)
            fc = self.fused_comm
            fc._ensure_gram_buffers(len(gram_specs))
            for slot, sp in enumerate(gram_specs):
                sub_idx = subsets[sp.index]
                if len(sub_idx) <= 1:
                    self.crafted_own[sp.index] = True
                    continue
                cen = center if center is not None else self.global_row
                fc.gram_attack(slot, sub_idx, sp.attack_mode, cen.contiguous(), a_cfg.search == "paper")
                info[f"gamma_{sp.index}"] = fc._last_gamma
                self.crafted_own[sp.index] = False
            torch.cuda.synchronize(self.device)
            self.ctx.barrier()                                  # every rank's slice has landed in every crafted row
            for slot, sp in enumerate(gram_specs):
                if not self.crafted_own.get(sp.index, False):
                    self.crafted[sp.index] = fc.craft_sym[slot].clone()
            return info
        for sp in attackers:
            if sp.attack_mode not in GRAM_MODES:
                continue
            sub = rows_all[subsets[sp.index]]
            if sub.shape[0] <= 1:
                self.crafted_own[sp.index] = True            # "return state_dict" (Utils.py:102-103)
                continue
            cen = center if center is not None else sub.mean(0)
            use_dev = native and a_cfg.metric == "flat"
            if use_dev:
                from .ops import flat, gram
                kk = sub.shape[0]
                mem = torch.ones(1, kk)
                _, mu, sd, _ = flat.fedavg(sub.contiguous(), torch.zeros(kk, device=self.device), member=mem,
                                           ref=cen.contiguous(), want_mu=True, want_sd=True)
                p = torch.sign(mu[0]) if sp.attack_mode == "Opt-Fang" else sd[0]
                a = torch.cat([sub - cen, (mu[0] - cen)[None], p[None]], dim=0)
                G = gram.gram(a, precision=a_cfg.gram_precision) if self.device.type == "cuda" else (a @ a.t())
                g4 = flat.gamma_search(G, kk, "sum" if sp.attack_mode == "Min-Sum" else "max", True,
                                       a_cfg.search == "paper")
                self.crafted[sp.index] = flat.craft_axpy(mu[0].contiguous(), p.contiguous(), g4,
                                                         sign_mode=False)
                info[f"gamma_{sp.index}"] = g4            # stays on the device; read lazily for metrics
            else:
                own = self.global_row if self.have_global else sub[0]
                row, meta = atk.craft(sp.attack_mode, own, sub, sp.attack_args, layout=self.layout, metric=a_cfg.metric,
                                      search=a_cfg.search, center=cen, gram_fn=self._gram_fn())
                self.crafted[sp.index] = row
                info[f"gamma_{sp.index}"] = meta.get("gamma")
            self.crafted_own[sp.index] = False
        return info

    def _apply_attacks(self, active: List[int]) -> None:
        for c in active:
            if c not in self.local:
                continue
            sp = self.specs[c]
            row = self._row_of(c)
            if sp.attack_mode == "Random":
                if not sp.attack_args:
                    raise TypeError("Random attack needs --attack_args <perturbation>")
                if self.backend == "native":
                    from .ops import flat
                    flat.craft_random(row, self.valid, float(sp.attack_args[0]),
                                      seed=int(self.cfg.server.random_seed or 0) * 1000003 + self.attempt,
                                      stream_id=c, out=row)
                else:
                    row.copy_(atk.random_attack(row, float(sp.attack_args[0]), self.valid, None))
            elif self.crafted_own.get(c, False):
                pass                                           # keeps the weights it received
            elif c in self.crafted:
                row.copy_(self.crafted[c])

    # ================================================================================== aggregation
    def _aggregate(self, rows_all: torch.Tensor, reporting: List[int]) -> Dict[str, object]:
        s = self.cfg.server
        info: Dict[str, object] = {}
        info_in, self._agg_info = getattr(self, "_agg_info", None), None      # deferred LIE request of this round
        native = self.backend == "native"
        sel = rows_all if len(reporting) == rows_all.shape[0] else rows_all[reporting]
        sizes = [self.sizes[c] for c in reporting]
        gram_fn = self._gram_fn()
        if self.mode == "fedavg":
            if native:
                from .ops import flat
                w = torch.tensor(sizes, dtype=torch.float32)
                w = (w / w.sum()).to(self.device)
                bc = self.rows[:self.CL] if (not self.ctx.distributed or self.fused_comm is None) else None
                lie_specs = info_in.pop("_lie_specs", None) if info_in is not None else None
                if lie_specs:
                    # K1 + K2 in ONE pass: FedAvg, broadcast and every LIE attacker's mean / std / crafted row
                    member = info_in.pop("_lie_member")[:, reporting]
                    _, _, _, lrows = flat.fedavg(sel.contiguous(), w, out=self.global_row, bcast=bc, member=member,
                                                 z=info_in.pop("_lie_z"), ref=info_in.pop("_lie_ref"), want_lie=True,
                                                 strict_k1=self.cfg.attack.strict_reference)
                    for q, sp in enumerate(lie_specs):
                        self.crafted[sp.index] = lrows[q]
                        self.crafted_own[sp.index] = False
                else:
                    flat.fedavg(sel.contiguous(), w, out=self.global_row, bcast=bc)
                self._rows_hold_global = bc is not None
            else:
                self.global_row.copy_(agg.fedavg(sel, sizes))
        elif self.mode == "trimmed_mean":
            k = int(len(reporting) * s.trim_ratio)
            if 2 * k >= len(reporting):
                raise ValueError("too few clients for the chosen trim_ratio")
            if native and len(reporting) <= 256:
                from .ops import flat
                flat.coord_select(sel.contiguous(), "trimmed", k, out=self.global_row)
            else:
                self.global_row.copy_(agg.trimmed_mean(sel, s.trim_ratio))
        elif self.mode == "median":
            if native and len(reporting) <= 256:
                from .ops import flat
                flat.coord_select(sel.contiguous(), "median", out=self.global_row)
            else:
                self.global_row.copy_(agg.median(sel))
        elif self.mode == "krum":
            kf = int(sel.shape[0] * s.genuine_rate) if s.krum_reference_f else s.krum_f
            i = agg.krum_select(sel, kf, gram_fn, reference_slice=s.krum_reference_f)
            info["krum_winner"] = reporting[i]
            self.global_row.copy_(sel[i])
        elif self.mode == "shieldfl":
            w = agg.shieldfl_weights(sel, gram_fn).to(self.device)
            if native:
                from .ops import flat
                flat.fedavg(sel.contiguous(), w, out=self.global_row)
            else:
                self.global_row.copy_(agg.weighted_sum(sel, w))
        elif self.mode == "scionfl":
            scores = None
            if self.backend == "native":                          # bit-packed pass: deq rows are never built
                from .ops import flat
                scores, _ = flat.scion_cosine_distances(sel.contiguous(), self.valid, seed=int(self.cfg.server.random_seed or 0) * 7919 + self.attempt)
                scores = scores.tolist()
            gen = None
            if scores is None:
                # rank-independent stream: every rank of a replicated aggregation must draw the same Bernoulli bits
                gen = torch.Generator(device=sel.device)
                gen.manual_seed(int(self.cfg.server.random_seed or 0) * 7919 + self.attempt)
            out, kept = agg.scionfl(sel, sizes, valid=self.valid.bool(), generator=gen, keep=s.scionfl_keep,
                                    scores=scores)
            info["scionfl_kept"] = [reporting[i] for i in kept]
            self.global_row.copy_(out)
        elif self.mode == "FLTrust":
            new, ts = agg.fltrust(sel, self._fltrust_base, self.server_row, gram_fn)
            info["trust"] = ts.tolist()
            self.global_row.copy_(new)
        elif self.mode == "gmm":
            kept = agg.gmm_filter(sel, [self.specs[c].attack for c in reporting])
            info["gmm_kept"] = [reporting[i] for i in kept]
            if not kept:
                info["failed"] = True
            else:
                self.global_row.copy_(sel[kept].mean(0))
        elif self.mode == "fltracer":
            bad = set(agg.fltracer_anomalies(sel))
            kept = [i for i in range(len(reporting)) if i not in bad] or list(range(len(reporting)))
            info["fltracer_removed"] = [reporting[i] for i in sorted(bad)]
            self.global_row.copy_(agg.fedavg(sel[kept], [sizes[i] for i in kept]))
        elif self.mode == "byzantine":
            self.global_row.copy_(agg.byzantine_tolerance(sel))
        else:
            raise ValueError(f"Server mode '{self.mode}' is not valid.")
        self.have_global = True
        return info

    def _aggregate_fused(self, reporting: List[int], center: Optional[torch.Tensor], info: Dict[str, object]) -> None:
        """FedAvg + LIE moments + broadcast in one kernel per rank over NVLink peer memory (csrc/comm.cu)."""
        w = torch.zeros(self.C, dtype=torch.float32)
        for c in reporting:
            w[c] = float(self.sizes[c])
        w = w / w.sum()
        lie_specs = info.pop("_lie_specs", None)
        member, z = info.pop("_lie_member", None), info.pop("_lie_z", None)
        info.pop("_lie_ref", None)
        ref = center if center is not None else self.init_row
        rows = self.fused_comm.fedavg(w, member, z, ref.contiguous(), self.cfg.attack.strict_reference)
        if lie_specs:
            for q, sp in enumerate(lie_specs):
                self.crafted[sp.index] = rows[q]
                self.crafted_own[sp.index] = False
        self.have_global = True
        self._rows_hold_global = True

    SLICED_MODES = ("median", "trimmed_mean", "krum", "shieldfl", "FLTrust", "scionfl")

    def _aggregate_sliced(self, reporting: List[int], info: Dict[str, object]) -> None:
        """Defences on N GPUs without gathering a row: rank r owns coordinate slice r.  Median / trimmed mean read the C
        values of their coordinates from the clients' rows in place (peer pointers) and store the result into every
        rank's global model (K6); Krum / ShieldFL / FLTrust need only an n x n Gram: every rank computes the Gram of its
        K-slice (tcgen05), the partials are summed by K1-fused, the weights are replicated scalars and the weighted
        aggregation is the K1-fused kernel (``/root/reference/src/Utils.py:267-302,326-357``; ``server.py:306-350,703-743``)."""
        from .ops import flat
        s, fc = self.cfg.server, self.fused_comm
        sl = self.D // fc.world
        lo, hi = fc.rank * sl, (fc.rank + 1) * sl
        ptrs = [fc.row_ptr(c) for c in reporting]
        n = len(reporting)
        if self.mode in ("median", "trimmed_mean"):
            k = int(n * s.trim_ratio) if self.mode == "trimmed_mean" else 0
            if self.mode == "trimmed_mean" and 2 * k >= n:
                raise ValueError("too few clients for the chosen trim_ratio")
            flat.coord_select_ptrs(ptrs, lo, hi, [int(p) for p in fc.h_glob.buffer_ptrs],
                                   "median" if self.mode == "median" else "trimmed", k, self.device)
            torch.cuda.synchronize(self.device)
            self.ctx.barrier()                                    # every owner's slice has landed in every global buffer
            self.have_global = True
            return
        prec = self.cfg.attack.gram_precision
        zero = getattr(self, "_zero_row", None)
        if zero is None:
            zero = self._zero_row = torch.zeros(self.D, device=self.device)
        w = torch.zeros(self.C, dtype=torch.float32)
        g0_weight = 0.0
        if self.mode == "krum":
            cen = self.global_row if self.have_global else zero
            G = flat.sliced_gram(ptrs, cen.contiguous(), lo, hi, prec, fc.allreduce_small)
            kf = int(n * s.genuine_rate) if s.krum_reference_f else s.krum_f
            i = agg.krum_select(torch.zeros(n, 1, device=self.device), kf, lambda _r: G, reference_slice=s.krum_reference_f)
            info["krum_winner"] = reporting[i]
            w[reporting[i]] = 1.0
        elif self.mode == "scionfl":
            # only the 1-bit codes cross NVLink (server.py:436-492); the kept clients are averaged by K1-fused
            cosd = fc.scion_distances(reporting, int(s.random_seed or 0) * 7919 + self.attempt)
            kept = agg.scionfl_keep(cosd.tolist(), s.scionfl_keep)
            info["scionfl_kept"] = [reporting[i] for i in kept]
            tot = float(sum(self.sizes[reporting[i]] for i in kept))
            for i in kept:
                w[reporting[i]] = float(self.sizes[reporting[i]]) / tot
        elif self.mode == "shieldfl":
            G = flat.sliced_gram(ptrs, zero, lo, hi, prec, fc.allreduce_small)      # cosines of the raw vectors: no centring
            a = agg.shieldfl_weights(torch.zeros(n, 1, device=self.device), lambda _r: G).cpu()
            for j, c in enumerate(reporting):
                w[c] = float(a[j])
        else:                                                    # FLTrust: deltas against the round's base model
            base = self._fltrust_base
            G = flat.sliced_gram(ptrs + [self.server_row.data_ptr()], base.contiguous(), lo, hi, prec,
                                 fc.allreduce_small).to(torch.float64)
            norms = torch.diagonal(G).clamp_min(0).sqrt()
            cos = G[:n, n] / torch.clamp(norms[:n] * norms[n], min=1e-8)
            ts = cos.clamp_min(0)
            scale = ((norms[n] / (norms[:n] + 1e-6)) * ts / (ts.sum() + 1e-6)).cpu()
            info["trust"] = ts.to(torch.float32).tolist()
            for j, c in enumerate(reporting):
                w[c] = float(scale[j])
            g0_weight = 1.0 - float(scale.sum())                  # new = g0 + sum a_i (w_i - g0)
        fc.fedavg(w, None, None, zero, False)
        if g0_weight != 0.0:
            self.global_row.add_(self._fltrust_base, alpha=g0_weight)
            self._rows_hold_global = False
        else:
            self._rows_hold_global = True
        self.have_global = True

    def _fltrust_server_update(self) -> None:
        """Server trains on the root set: first 200 test samples, batch 100, no shuffle (server.py:289-294,711)."""
        s, l = self.cfg.server, self.cfg.learning
        n = min(s.fltrust_root_size, self.test_dev.shape[0])
        base = self.global_row if self.have_global else self.init_row
        self._fltrust_base = base.clone()
        self.server_row.copy_(base)
        idx = torch.arange(n, dtype=torch.int32, device=self.device)[None, None, :].expand(1, l.epoch, n).contiguous()
        if self.backend == "native":
            T = self.T
            plan = T.make_plan(self.train_prog, s.fltrust_root_batch, 1, self.D, self.device)
            plan.chunk = 1
            gp = torch.zeros(plan.chunk * plan.G, self.D, device=self.device)
            w = self.server_row[None].contiguous()
            m, v = torch.zeros_like(w), torch.zeros_like(w)
            T.train_round(self.train_prog, plan, w, m, v, gp, self.test_dev, idx,
                          torch.tensor([n], dtype=torch.int32, device=self.device),
                          torch.tensor([self.C], dtype=torch.int32, device=self.device), s.fltrust_root_batch,
                          l.epoch, optimizer=l.optimizer, lr=l.learning_rate, momentum=l.momentum,
                          seed=self.attempt + 77, dropout=l.dropout)
            self.server_row.copy_(w[0])
        else:
            keep_rows, keep_m, keep_v, keep_cl, keep_data = self.rows, self.m, self.v, self.CL, self.train_dev
            self.rows, self.m, self.v, self.CL = self.server_row[None].clone(), torch.zeros(1, self.D, device=self.device), torch.zeros(1, self.D, device=self.device), 1
            self.train_dev = self.test_dev
            gen_state = self.dev_gen.get_state()
            bs = l.batch_size
            l.batch_size = s.fltrust_root_batch
            try:
                self._train_torch([n])
                self.server_row.copy_(self.rows[0])
            finally:
                l.batch_size = bs
                self.rows, self.m, self.v, self.CL, self.train_dev = keep_rows, keep_m, keep_v, keep_cl, keep_data
                self.dev_gen.set_state(gen_state)

    # ================================================================================== hyper
    def _hyper_round(self, rows_all: torch.Tensor, reporting: List[int]) -> Dict[str, object]:
        s, l = self.cfg.server, self.cfg.learning
        info: Dict[str, object] = {}
        if self.native_hyper is not None:
            self.native_hyper.unpack()
        snap = self.hyper.snapshot()          # detection rollback (server.py:297-298) and failed-round rollback (Q10)
        self._hyper_snap = snap
        order = [c for c in self.selected if c in reporting]
        if self._hyper_sharded():
            # multi-GPU: row-sharded head, one persistent kernel per rank; client rows are read in place over NVLink and
            # the personalised models are scattered back into the owners' rows (no all-gather)
            nh = self.native_hyper
            nh.round(order, [self.fused_comm.row_ptr(c) for c in order], l.hyper_lr, l.clip_grad_norm)
            nh.emit_rows(reporting, [self.fused_comm.row_ptr(c) for c in reporting])      # server.py:674-678
            torch.cuda.synchronize(self.device)
            if nh.check_round_error():
                info["failed"] = True
            self.ctx.barrier()
            self._rows_hold_theta = True
        elif self.native_hyper is not None and s.hyper_update == "sequential" and self.device.type == "cuda":
            # one persistent launch for all clients of the round (4 launches per client before)
            self.native_hyper.round(order, [rows_all[c].data_ptr() for c in order], l.hyper_lr, l.clip_grad_norm)
        elif self.native_hyper is not None and s.hyper_update == "sequential":
            for c in order:
                self.native_hyper.step(c, rows_all[c].contiguous(), l.hyper_lr, l.clip_grad_norm)
        else:
            if self.native_hyper is not None:
                self.native_hyper.unpack()
            agg.hyper_round_torch(self.hyper, order, rows_all[order], l.hyper_lr, l.clip_grad_norm, s.hyper_update)
            if self.native_hyper is not None:
                self.native_hyper.pack()
        if self.monitor is not None:
            if self.native_hyper is not None:
                self.native_hyper.unpack()
            emb = {c: self.hyper.embedding(c).detach().cpu().numpy() for c in self.selected}
            remove = self.monitor.observe(self.num_round - self.remaining + 1, list(self.selected), emb)
            if remove:
                for c in remove:
                    print(f"Removing anomaly {c}, rolling back")
                    self.selected.remove(c)
                self.hyper.restore(snap)                          # server.py:536 (parameters only)
                if self.native_hyper is not None:
                    self.native_hyper.pack()
                info["removed"] = remove
        return info

    # ================================================================================== validation
    def _validate(self, reporting: List[int]) -> bool:
        if self.data_name == "ICU":
            if self.mode == "hyper" and self._hyper_sharded():
                # personalised models already sit in their owners' rows: every rank scores its own clients, the
                # probabilities (n_test floats per client) are gathered and pooled into one ROC-AUC like test_hyper_icu
                n_rep = len(reporting)
                mine = [j for j, c in enumerate(self.local) if c < n_rep]
                cpg = self.ctx.clients_per_rank(self.C)
                probs_loc = torch.zeros(cpg, self.test_dev.shape[0], device=self.device)
                if mine:
                    probs_loc[:len(mine)] = self.T.infer(self.infer_prog, self.rows[mine].contiguous(), self.test_dev)
                probs_all = self.ctx.all_gather_rows(probs_loc, self.C)[:n_rep]
                labels = self.test_labels[None].expand(probs_all.shape[0], -1)
                return self.validator.score_icu(probs_all.reshape(-1), labels.reshape(-1))
            if self.mode == "hyper":
                ids = list(range(len(reporting)))                 # test_hyper(hnet, len(all_model_parameters))
                if self.native_hyper is not None:
                    rows = self.native_hyper.emit(ids)
                else:
                    rows = self.hyper.emit_torch(torch.tensor(ids, device=self.device))
            else:
                rows = self.global_row[None]
            if self.backend == "native":
                probs = self.T.infer(self.infer_prog, rows.contiguous(), self.test_dev)
            else:
                mod = self.module.to(self.device).eval()
                outs = []
                with torch.no_grad():
                    for r in rows:
                        outs.append(torch.func.functional_call(mod, dict(self.layout.views(r)),
                                                               (self.test_dev[:, :7], self.test_dev[:, 7:23]))[:, 0])
                probs = torch.stack(outs)
            labels = self.test_labels[None].expand(probs.shape[0], -1)
            return self.validator.score_icu(probs.reshape(-1), labels.reshape(-1))
        if self.har_trainer is not None:
            tx = self.test_x.reshape(self.test_x.shape[0], -1)
            mb = self.har_trainer.max_batch
            logits = torch.cat([self.har_trainer.logits_of(self.global_row, tx[i:i + mb]).clone()
                                for i in range(0, tx.shape[0], mb)])
            return self.validator.score_har(logits, self.test_y)
        mod = self.module.to(self.device).eval()
        with torch.no_grad():
            logits = torch.cat([torch.func.functional_call(mod, dict(self.layout.views(self.global_row)), (self.test_x[i:i + 256],))
                                for i in range(0, self.test_x.shape[0], 256)])
        return self.validator.score_har(logits, self.test_y)

    # ================================================================================== one round
    def run_round(self) -> RoundResult:
        """One START -> train -> UPDATE -> aggregate -> validate cycle (``server.py:205-275,277-567``).

        Observability / robustness hooks around it: an NVTX range per phase, an optional
        ``cudaProfilerStart/Stop`` bracket on attempt ``runtime.profile-round`` (so ``ncu --profile-from-start off``
        captures exactly one round), the rank-health watchdog, and the test-only fault injectors."""
        self.attempt += 1
        rt = self.cfg.runtime
        profiling = rt.profile_round > 0 and self.attempt == rt.profile_round and self.device.type == "cuda"
        if profiling:
            torch.cuda.synchronize(self.device)
            torch.cuda.cudart().cudaProfilerStart()
        self.watchdog.arm(f"round attempt {self.attempt}")
        try:
            return self._run_round()
        finally:
            self.watchdog.disarm()
            if profiling:
                torch.cuda.synchronize(self.device)
                torch.cuda.cudart().cudaProfilerStop()

    def _faults(self, kind: str) -> List[int]:
        """Clients hit by a ``runtime.fault-inject`` entry of ``kind`` on this attempt (test-only hook)."""
        return [int(f["client"]) for f in self.cfg.runtime.fault_inject
                if f.get("kind") == kind and int(f.get("attempt", -1)) == self.attempt]

    def _run_round(self) -> RoundResult:
        s, l = self.cfg.server, self.cfg.learning
        round_no = self.num_round - self.remaining + 1
        t_wall = time.perf_counter()
        # The rollback snapshot belongs to THIS attempt: a round that fails before aggregation must not restore the
        # snapshot of an earlier, successful round (that would silently undo a counted and checkpointed update).
        self._hyper_snap = None
        trace = [] if os.environ.get("ATTACKFL_HOST_TRACE") else None     # host timeline of the round (debug aid)

        def mark(name):
            if trace is not None:
                trace.append((name, round((time.perf_counter() - t_wall) * 1e3, 3)))
        for _ in self._faults("stall"):                           # exercises the watchdog
            time.sleep(max(self.cfg.runtime.round_timeout_s * 2, 0.2))
        if self.host_train is not None:
            with self.timer.phase("h2d"):
                self.train_dev.copy_(self.host_train, non_blocking=True)
        starting = list(self.selected)
        # ---- START (notify_clients, server.py:569-614)
        with self.timer.phase("distribute"):
            self._distribute(starting)
        active_attackers: List[int] = []
        for c in starting:
            self.training_round[c] += 1
            self.sizes[c] = self.client_rng[c].randrange(s.num_data_range[0], s.num_data_range[1] + 1)
            sp = self.specs[c]
            has_models = (c in self.crafted) or self.crafted_own.get(c, False) or \
                         (sp.attack_mode == "Random" and self.genuine_prev)
            if sp.attack and self.training_round[c] >= sp.attack_round and has_models:
                active_attackers.append(c)
        if self.mode == "FLTrust":
            with self.timer.phase("fltrust_root"):
                self._fltrust_server_update()
        # ---- local work
        with self.timer.phase("craft_apply"):
            self._apply_attacks(active_attackers)
        sizes_local = [0 if (c in active_attackers or c not in starting) else self.sizes[c] for c in self.local]
        with self.timer.phase("train"):
            if self.CL > 0:
                if self.har_trainer is not None:
                    loss = self._train_har_native(sizes_local)
                else:
                    loss = self._train_native(sizes_local) if self.backend == "native" else self._train_torch(sizes_local)
            else:
                loss = torch.zeros(1, 2, device=self.device)
        for c in self._faults("nan-loss"):                        # client c reports result=False (client.py:100-102)
            if c in self.local:
                loss[self.local.index(c), 1] = 1.0
        for c in self._faults("nan-weights"):                     # a diverged update: caught by the NaN validation
            if c in self.local:
                self.rows[self.local.index(c)].fill_(float("nan"))
        # ---- UPDATE: result flags + losses (the only per-client scalars that reach the host)
        mark("train_enqueued")
        flags = torch.zeros(self.C, 3, device=self.device)
        if self.local:                                            # local clients are one contiguous block of ids
            lo_c, n_c = self.local[0], len(self.local)
            flags[lo_c:lo_c + n_c, 0] = loss[:n_c, 1]
            flags[lo_c:lo_c + n_c, 1] = loss[:n_c, 0]
            flags[lo_c:lo_c + n_c, 2] = torch.tensor([float(max(x, 0)) for x in sizes_local], device=self.device)
        if self.ctx.distributed:
            self.ctx.all_reduce_sum(flags)
        reporting = list(starting)
        info: Dict[str, object] = {}
        fused = self.fused_comm is not None and self.mode == "fedavg"
        gram_sliced = fused and self.cfg.attack.metric == "flat" and self.have_global
        need_gather = (not fused) or (any(sp.attack and sp.attack_mode in GRAM_MODES for sp in self.specs)
                                      and not gram_sliced)
        if self._hyper_sharded() and not any(sp.attack for sp in self.specs):
            need_gather = False                                   # rows are read in place by the sharded hyper kernel
        sliced = (self.fused_comm is not None and self.backend == "native" and self.mode in self.SLICED_MODES
                  and len(starting) <= 256 and self.cfg.attack.metric == "flat")
        has_lie = any(sp.attack and sp.attack_mode == "LIE" for sp in self.specs)
        if sliced and not has_lie and (self.have_global or not any(sp.attack and sp.attack_mode in GRAM_MODES for sp in self.specs)):
            need_gather = False                                   # slice-owned defences read the rows in place
            gram_sliced = self.have_global
        rows_all = None
        if need_gather:
            with self.timer.phase("gather"):
                rows_all = self._all_rows()
        mark("flags_enqueued")
        flags_h = flags.cpu()                                     # D2H: 3 floats per client
        mark("flags_on_host")
        ok = bool((flags_h[:, 0] == 0).all()) and bool(torch.isfinite(flags_h[:, 1]).all())
        losses = [float(flags_h[c, 1] / max(float(flags_h[c, 2]) * l.epoch, 1.0)) for c in range(self.C)]
        prev_global = self.global_row.clone() if self.have_global else None
        if ok:
            genuine = [c for c in reporting if not self.specs[c].attack]
            center = self.global_row.clone() if (self.have_global and self.mode != "hyper") else None
            with self.timer.phase("craft"):
                one_pass = (self.backend == "native" and self.mode == "fedavg" and not self.ctx.distributed
                            and rows_all is not None)
                info.update(self._prepare_crafts(rows_all, genuine, center, defer_lie=fused or one_pass))
            self.genuine_prev = genuine
            if self.fused_comm is not None and os.environ.get("ATTACKFL_DEBUG_OWNERSHIP"):
                self.fused_comm.assert_ownership(reporting)
            mark("crafted")
            with self.timer.phase("aggregate"):
                if sliced and rows_all is None:
                    self._aggregate_sliced(reporting, info)
                    err = self.fused_comm.check_error(raise_on_error=False)
                    if err:
                        self._log(f"fused aggregation reported error {err}; the round is discarded", "red")
                        self.fused_comm.reset_error()
                        info["failed"] = True
                elif fused:
                    self._aggregate_fused(reporting, center, info)
                    # a peer-barrier time-out leaves glob / lie / the client rows stale: fail the round on every rank
                    # (the error word is part of the symmetric control block, every rank reads its own copy)
                    err = self.fused_comm.check_error(raise_on_error=False)
                    if err:
                        self._log(f"fused aggregation reported error {err}; the round is discarded", "red")
                        self.fused_comm.reset_error()
                        if prev_global is not None:
                            self.global_row.copy_(prev_global)
                        info["failed"] = True
                elif self.mode == "hyper":
                    info.update(self._hyper_round(rows_all, reporting))
                else:
                    self._agg_info = info
                    info.update(self._aggregate(rows_all, reporting))
            if info.get("failed"):
                ok = False
            mark("aggregated")
        metric = None
        if ok and s.validation:
            with self.timer.phase("validate"):
                ok = self._validate(reporting)
                metric = self.validator.last_metric
        mark("validated")
        if trace is not None:
            info["host_trace_ms"] = trace
        # ---- decision (server.py:545-567)
        if not ok:
            self._log("Training failed!", "yellow")
            self.failures_in_a_row += 1
            if prev_global is not None and not self.cfg.attack.strict_reference:
                self.global_row.copy_(prev_global)               # do not propagate a poisoned global (Q10)
                self._rows_hold_global = False
            if self.mode == "hyper" and getattr(self, "_hyper_snap", None) is not None and not self.cfg.attack.strict_reference:
                self.hyper.restore(self._hyper_snap, optimizer_state=True)
                if self.native_hyper is not None:
                    self.native_hyper.pack()
        else:
            self.failures_in_a_row = 0
            head_full = None
            if self._hyper_sharded() and self.fused_comm.any_rank_saves:
                head_full = self.fused_comm.hyper_gather_full()  # every rank joins; rank 0 then writes the .pth
            if self.save and self.ctx.rank == 0:
                self.save_checkpoint(head_full)
            self.remaining -= 1
        timings = self.timer.collect()
        timings["wall"] = (time.perf_counter() - t_wall) * 1e3
        res = RoundResult(round_no, ok, metric, losses, list(self.sizes), active_attackers, timings, info)
        self.history.append(res)
        self.metrics.write({"round": round_no, "attempt": self.attempt, "ok": ok, "metric": metric,
                            "attackers": active_attackers, "timings_ms": timings,
                            "mean_loss": float(np.mean([x for x in losses if x == x])) if losses else None})
        return res

    # ================================================================================== driver loop
    def run(self, max_rounds: Optional[int] = None) -> List[RoundResult]:
        out = []
        while self.remaining > 0:
            self._log(f"Start training round {self.num_round - self.remaining + 1}", "yellow")
            res = self.run_round()
            out.append(res)
            mr = self.cfg.server.max_retries
            if not res.ok and mr and self.failures_in_a_row > mr:
                raise RuntimeError(f"round {res.round} failed {self.failures_in_a_row} times (server.max-retries={mr})")
            if max_rounds is not None and len(out) >= max_rounds:
                break
        return out

    # ================================================================================== checkpoint
    def save_checkpoint(self, head_full: Optional[Dict[str, torch.Tensor]] = None) -> None:
        """Writes ``<model>.pth`` / ``<model>_hyper_<C>.pth``.  With a row-sharded hyper-network ``head_full`` is the
        gathered head (``FusedComm.hyper_gather_full``, a collective the caller runs on every rank first)."""
        import json
        if self.mode == "hyper":
            if self.native_hyper is not None:
                self.native_hyper.unpack()
            path = os.path.join(self.workdir, f"{self.model_name}_hyper_{self.C}.pth")
            torch.save(self.hyper.export_state_dict(head_full=head_full), path + ".tmp")
            os.replace(path + ".tmp", path)
        else:
            self.layout.save_pth(self.global_row, os.path.join(self.workdir, f"{self.model_name}.pth"))
        side = {"round_done": self.num_round - self.remaining + 1, "attempt": self.attempt,
                "config": self.cfg.digest(), "seed": self.cfg.server.random_seed, "mode": self.mode}
        with open(os.path.join(self.workdir, f"{self.model_name}.ckpt.json"), "w") as fh:
            json.dump(side, fh)

    def state_dict(self) -> Dict[str, torch.Tensor]:
        """Current global model (or, in hyper mode, the hyper-network) with the reference's keys."""
        if self.mode == "hyper":
            head_full = None
            if self._hyper_sharded():                            # a collective: call on every rank
                head_full = self.fused_comm.hyper_gather_full(all_ranks=True)
            if self.native_hyper is not None:
                self.native_hyper.unpack()
            return self.hyper.export_state_dict(head_full=head_full)
        return self.layout.unflatten(self.global_row, device="cpu")

    def close(self) -> None:
        self.metrics.close()
