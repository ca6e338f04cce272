"""Hyper-networks: reference-structured ``HyperNetwork`` / ``CNNHyper`` and the fused ``FusedHyperNet``.

Reference parity: ``/root/reference/src/Model.py:251-304`` (``HyperNetwork``: Embedding(n,8) ->
MLP 8-100-100-100 -> one ``Linear(100, numel)`` head per ``state_dict`` entry of the target) and
``:309-416`` (``CNNHyper``, dead code in the reference, kept for API completeness).

``FusedHyperNet`` is the engine's representation: the per-tensor heads are
concatenated in flat-layout order into ``W_h [D_pad, H]`` and ``b_h [D_pad]`` so that emitting a
client model is one GEMV/GEMM row-slice (kernel K5a) and the pFedHN update is a rank-1 step
(kernel K5b).  ``export_state_dict`` / ``import_state_dict`` split / merge to the reference's
``{model}_hyper_{N}.pth`` key layout (``target_model.*``, ``embeddings.weight``,
``mlp.{0,2,4}.*``, ``hyper_layers.<key with '.'->'__'>.{weight,bias}``; ``server.py:551``).
"""
from __future__ import annotations

from collections import OrderedDict
from typing import Dict, Mapping, Optional, Tuple

import torch
import torch.nn as nn
from torch.nn.utils import spectral_norm

from ..layout import FlatLayout


def _maybe_sn(layer: nn.Module, on: bool) -> nn.Module:
    return spectral_norm(layer) if on else layer


def _make_mlp(embedding_dim: int, hidden_dim: int, n_hidden: int, spec_norm: bool) -> nn.Sequential:
    mods = [_maybe_sn(nn.Linear(embedding_dim, hidden_dim), spec_norm)]
    for _ in range(n_hidden):
        mods += [nn.ReLU(inplace=True), _maybe_sn(nn.Linear(hidden_dim, hidden_dim), spec_norm)]
    return nn.Sequential(*mods)


class HyperNetwork(nn.Module):
    """Generic hyper-network; ``forward(idx) -> (OrderedDict name->tensor, embedding)``.

    Like the reference it registers ``target_model`` as a sub-module, so ``parameters()`` and
    ``state_dict()`` contain ``target_model.*`` entries which never receive gradients
    (SURVEY §2.9 (iv)).
    """

    def __init__(self, target_model: nn.Module, n_nodes: int, embedding_dim: int,
                 hidden_dim: int = 100, spec_norm: bool = False, n_hidden: int = 2):
        super().__init__()
        self.target_model = target_model
        self.embeddings = nn.Embedding(num_embeddings=n_nodes, embedding_dim=embedding_dim)
        self.mlp = _make_mlp(embedding_dim, hidden_dim, n_hidden, spec_norm)
        self._shapes = OrderedDict((k, tuple(v.shape)) for k, v in target_model.state_dict().items())
        self.hyper_layers = nn.ModuleDict({
            k.replace(".", "__"): _maybe_sn(nn.Linear(hidden_dim, int(torch.Size(s).numel())), spec_norm)
            for k, s in self._shapes.items()})

    def forward(self, idx: torch.Tensor):
        emd = self.embeddings(idx)
        feat = self.mlp(emd)
        out = OrderedDict()
        for safe, head in self.hyper_layers.items():
            name = safe.replace("__", ".")
            w = head(feat)
            out[name] = w.view(-1) if "bias" in name else w.view(self._shapes[name])
        return out, emd


class CNNHyper(nn.Module):
    """Hand-specialised hyper-network for ``CNNModel`` (reference: dead code, Model.py:309-416)."""

    _TARGET = (
        ("vitals_conv1", (32, 1, 3)), ("vitals_conv2", (64, 32, 3)), ("vitals_conv3", (128, 64, 3)),
        ("labs_conv1", (32, 1, 3)), ("labs_conv2", (64, 32, 3)), ("labs_conv3", (128, 64, 3)),
        ("fc1", (128, 1024)), ("fc2", (64, 128)), ("fc3", (32, 64)), ("output", (1, 32)),
    )

    def __init__(self, n_nodes: int, embedding_dim: int, hidden_dim: int, n_hidden: int,
                 spec_norm: bool = False):
        super().__init__()
        self.embeddings = nn.Embedding(num_embeddings=n_nodes, embedding_dim=embedding_dim)
        self.mlp = _make_mlp(embedding_dim, hidden_dim, n_hidden, spec_norm)
        for name, shape in self._TARGET:
            numel = int(torch.Size(shape).numel())
            setattr(self, f"{name}_weights", _maybe_sn(nn.Linear(hidden_dim, numel), spec_norm))
            setattr(self, f"{name}_bias", _maybe_sn(nn.Linear(hidden_dim, shape[0]), spec_norm))

    def forward(self, idx: torch.Tensor):
        emd = self.embeddings(idx)
        feat = self.mlp(emd)
        out = OrderedDict()
        for name, shape in self._TARGET:
            out[f"{name}.weight"] = getattr(self, f"{name}_weights")(feat).view(shape)
            out[f"{name}.bias"] = getattr(self, f"{name}_bias")(feat).view(-1)
        return out, emd


class FusedHyperNet:
    """Engine-side hyper-network with the heads fused into ``W_h [D_pad, H]`` / ``b_h [D_pad]``.

    Tensors (all fp32, on ``device``):
      ``emb [N, E]``, ``mlp_w[i]``, ``mlp_b[i]`` (3 layers: E->H, H->H, H->H), ``head_w [D_pad, H]``,
      ``head_b [D_pad]``, ``target [D_pad]`` (the passive ``target_model.*`` entries, exported only).
    Adam state mirrors every trainable tensor (``m_*``, ``v_*``) and a shared step counter, matching
    ``torch.optim.Adam(hnet.parameters(), lr=hyper_lr)`` (server.py:165).
    """

    def __init__(self, layout: FlatLayout, n_nodes: int, embedding_dim: int = 8, hidden_dim: int = 100,
                 n_hidden: int = 2, device="cpu"):
        self.layout = layout
        self.n_nodes, self.E, self.H, self.n_hidden = n_nodes, embedding_dim, hidden_dim, n_hidden
        self.device = torch.device(device)
        z = lambda *s: torch.zeros(*s, dtype=torch.float32, device=self.device)
        self.emb = z(n_nodes, embedding_dim)
        dims = [embedding_dim] + [hidden_dim] * (n_hidden + 1)
        self.mlp_w = [z(dims[i + 1], dims[i]) for i in range(n_hidden + 1)]
        self.mlp_b = [z(dims[i + 1]) for i in range(n_hidden + 1)]
        self.head_w = z(layout.padded, hidden_dim)
        self.head_b = z(layout.padded)
        self.target = z(layout.padded)
        self.step = 0
        self.row_lo, self.row_hi = 0, layout.padded         # rows of the head held by this object (all of them by default)
        self._init_adam()

    # -- parameters in torch's ``hnet.parameters()`` order minus the grad-less ``target_model.*``
    def trainable(self) -> "OrderedDict[str, torch.Tensor]":
        d = OrderedDict()
        d["emb"] = self.emb
        for i in range(len(self.mlp_w)):
            d[f"mlp_w{i}"] = self.mlp_w[i]
            d[f"mlp_b{i}"] = self.mlp_b[i]
        d["head_w"] = self.head_w
        d["head_b"] = self.head_b
        return d

    def _init_adam(self):
        self.m = OrderedDict((k, torch.zeros_like(v)) for k, v in self.trainable().items())
        self.v = OrderedDict((k, torch.zeros_like(v)) for k, v in self.trainable().items())
        self.step = 0

    # ------------------------------------------------------------------ row sharding (multi-GPU, SURVEY 2.11 ZeRO row)
    @property
    def sharded(self) -> bool:
        return self.row_lo != 0 or self.row_hi != self.layout.padded

    def shard_rows(self, lo: int, hi: int) -> None:
        """Keep only rows ``[lo, hi)`` of ``W_h`` / ``b_h`` and of their Adam moments; the rest of the storage is
        released.  From then on this object holds ``1 / world`` of the head (the MLP / embedding stay replicated) and the
        full head only exists transiently when a checkpoint gathers it (``export_state_dict(head_full=...)``)."""
        assert not self.sharded and 0 <= lo < hi <= self.layout.padded
        for name in ("head_w", "head_b"):
            setattr(self, name, getattr(self, name)[lo:hi].clone())
            self.m[name] = self.m[name][lo:hi].clone()
            self.v[name] = self.v[name][lo:hi].clone()
        self.row_lo, self.row_hi = int(lo), int(hi)

    def _need_full(self, what: str) -> None:
        if self.sharded:
            raise RuntimeError(f"{what} needs the whole hyper-network head, but this rank holds rows "
                               f"[{self.row_lo}, {self.row_hi}) only")

    # ------------------------------------------------------------------ conversion
    @classmethod
    def from_module(cls, hnet: HyperNetwork, layout: Optional[FlatLayout] = None, device="cpu") -> "FusedHyperNet":
        layout = layout or FlatLayout.from_state_dict(hnet.target_model.state_dict())
        lin = [m for m in hnet.mlp if isinstance(m, nn.Linear)]
        self = cls(layout, hnet.embeddings.num_embeddings, hnet.embeddings.embedding_dim,
                   lin[0].out_features, len(lin) - 1, device)
        self.import_state_dict(hnet.state_dict())
        return self

    def import_state_dict(self, sd: Mapping[str, torch.Tensor]) -> None:
        dev = self.device
        self.emb.copy_(sd["embeddings.weight"].to(dev))
        for i in range(len(self.mlp_w)):
            self.mlp_w[i].copy_(sd[f"mlp.{2 * i}.weight"].to(dev))
            self.mlp_b[i].copy_(sd[f"mlp.{2 * i}.bias"].to(dev))
        self.head_w.zero_()
        self.head_b.zero_()
        self.target.zero_()
        for e in self.layout.entries:
            safe = e.name.replace(".", "__")
            a, b = max(e.offset, self.row_lo), min(e.offset + e.numel, self.row_hi)    # part of the entry held here
            if a < b:
                self.head_w[a - self.row_lo:b - self.row_lo].copy_(
                    sd[f"hyper_layers.{safe}.weight"][a - e.offset:b - e.offset].to(dev))
                self.head_b[a - self.row_lo:b - self.row_lo].copy_(
                    sd[f"hyper_layers.{safe}.bias"][a - e.offset:b - e.offset].to(dev))
            tk = f"target_model.{e.name}"
            if tk in sd:
                self.target[e.offset:e.offset + e.numel].copy_(sd[tk].reshape(-1).to(dev))

    def export_state_dict(self, device="cpu", head_full: Optional[Mapping[str, torch.Tensor]] = None
                          ) -> "OrderedDict[str, torch.Tensor]":
        """Same keys and order as ``HyperNetwork(...).state_dict()`` of the reference.  A row-sharded object needs
        ``head_full = {"head_w": [D_pad, H], "head_b": [D_pad]}`` (the shards gathered by the caller)."""
        if head_full is None:
            self._need_full("export_state_dict")
            head_w, head_b = self.head_w, self.head_b
        else:
            head_w, head_b = head_full["head_w"], head_full["head_b"]
        out = OrderedDict()
        for e in self.layout.entries:
            out[f"target_model.{e.name}"] = self.target[e.offset:e.offset + e.numel].view(e.shape).to(device).clone()
        out["embeddings.weight"] = self.emb.to(device).clone()
        for i in range(len(self.mlp_w)):
            out[f"mlp.{2 * i}.weight"] = self.mlp_w[i].to(device).clone()
            out[f"mlp.{2 * i}.bias"] = self.mlp_b[i].to(device).clone()
        for e in self.layout.entries:
            safe = e.name.replace(".", "__")
            out[f"hyper_layers.{safe}.weight"] = head_w[e.offset:e.offset + e.numel].to(device).clone()
            out[f"hyper_layers.{safe}.bias"] = head_b[e.offset:e.offset + e.numel].to(device).clone()
        return out

    def to_module(self, target_model: nn.Module) -> HyperNetwork:
        h = HyperNetwork(target_model, self.n_nodes, self.E, self.H, False, self.n_hidden)
        h.load_state_dict(self.export_state_dict())
        return h

    # ------------------------------------------------------------------ snapshot (detection rollback, Q10)
    def snapshot(self) -> Dict[str, object]:
        snap = {k: v.clone() for k, v in self.trainable().items()}
        return {"p": snap, "m": {k: v.clone() for k, v in self.m.items()},
                "v": {k: v.clone() for k, v in self.v.items()}, "step": self.step,
                "target": self.target.clone()}

    def restore(self, snap: Dict[str, object], optimizer_state: bool = False) -> None:
        """Reference rollback (server.py:536) restores parameters only; Adam moments keep running."""
        for k, v in self.trainable().items():
            v.copy_(snap["p"][k])
        self.target.copy_(snap["target"])
        if optimizer_state:
            for k in self.m:
                self.m[k].copy_(snap["m"][k])
                self.v[k].copy_(snap["v"][k])
            self.step = int(snap["step"])

    # ------------------------------------------------------------------ math (pure torch; oracle + CPU path)
    def features(self, idx: torch.Tensor) -> Tuple[torch.Tensor, list]:
        """MLP forward for client indices ``idx [n]`` -> (f ``[n,H]``, saved activations)."""
        e = self.emb[idx]
        acts = [e]
        h = e
        for i in range(len(self.mlp_w)):
            h = h @ self.mlp_w[i].t() + self.mlp_b[i]
            if i < len(self.mlp_w) - 1:
                acts.append(h)          # pre-activation
                h = torch.relu(h)
                acts.append(h)
        return h, acts

    def embedding(self, i: int) -> torch.Tensor:
        return self.emb[i:i + 1].clone()

    def emit_torch(self, idx: torch.Tensor) -> torch.Tensor:
        """``theta[n, D_pad] = f @ W_h^T + b_h`` (padding rows of W_h/b_h are zero)."""
        self._need_full("emit_torch")
        f, _ = self.features(idx)
        return f @ self.head_w.t() + self.head_b
