"""pFedHN-style hyper-network round on the fused representation (pure torch: oracle + CPU path).

Reference parity: ``Server.train_hyper`` (``/root/reference/server.py:637-678``): for each selected
client *sequentially*: ``theta_i = h(i)``; ``delta = theta_i - theta_i^client``;
``grad = (d theta_i / d phi)^T delta`` via ``autograd.grad``; ``clip_grad_norm_`` (effective here);
``Adam.step``; afterwards every reporting client's weights are regenerated from the updated
hyper-network (``:674-678``).

With the heads fused (``theta = W_h f + b_h``) the gradients are closed-form:
``dW_h = delta f^T`` (rank-1), ``db_h = delta``, ``df = W_h^T delta``, then a 3-layer MLP backward
and a one-row embedding gradient.  torch's Adam is *dense*: every trainable tensor (the whole
embedding table included) moves each step because of its running moments, and the grad-less
``target_model.*`` entries are skipped — both reproduced here.  The clip norm is
``sqrt(|delta|^2 (|f|^2 + 1) + |grad mlp|^2 + |grad emb|^2)``.

The native kernels K5a/K5b (``csrc/hyper.cu``) implement exactly this arithmetic on row-slices.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence

import torch

from ..models.hyper import FusedHyperNet

BETA1, BETA2, EPS = 0.9, 0.999, 1e-8


def _adam_inplace(p: torch.Tensor, g: torch.Tensor, m: torch.Tensor, v: torch.Tensor, lr: float, t: int) -> None:
    m.mul_(BETA1).add_(g, alpha=1 - BETA1)
    v.mul_(BETA2).addcmul_(g, g, value=1 - BETA2)
    bc1 = 1 - BETA1 ** t
    bc2 = 1 - BETA2 ** t
    denom = (v.sqrt() / math.sqrt(bc2)).add_(EPS)
    p.addcdiv_(m, denom, value=-lr / bc1)


def hyper_grads_torch(h: FusedHyperNet, client: int, theta_client: torch.Tensor) -> Dict[str, torch.Tensor]:
    h._need_full("hyper_grads_torch")
    idx = torch.tensor([client], device=h.device)
    f, acts = h.features(idx)                    # f [1,H]
    f = f[0]
    theta = h.head_w @ f + h.head_b
    delta = theta - theta_client
    grads: Dict[str, torch.Tensor] = {}
    grads["head_b"] = delta
    grads["head_w"] = torch.outer(delta, f)
    g = h.head_w.t() @ delta                     # dL/df  [H]
    # MLP backward; acts = [e, z0, a0, z1, a1] for 3 linear layers
    n = len(h.mlp_w)
    for i in reversed(range(n)):
        x_in = acts[0][0] if i == 0 else acts[2 * i][0]
        grads[f"mlp_w{i}"] = torch.outer(g, x_in)
        grads[f"mlp_b{i}"] = g
        g = h.mlp_w[i].t() @ g
        if i > 0:
            g = g * (acts[2 * i - 1][0] > 0).to(g.dtype)
    ge = torch.zeros_like(h.emb)
    ge[client] = g
    grads["emb"] = ge
    grads["_delta_sq"] = (delta * delta).sum()
    return grads


def hyper_step_torch(h: FusedHyperNet, client: int, theta_client: torch.Tensor, lr: float,
                     clip: float) -> float:
    """One sequential pFedHN step for ``client``; returns the pre-clip gradient norm."""
    grads = hyper_grads_torch(h, client, theta_client)
    grads.pop("_delta_sq")
    total = math.sqrt(sum(float((g.double() ** 2).sum()) for g in grads.values()))
    if clip and clip > 0:
        coef = min(1.0, clip / (total + 1e-6))
        if coef < 1.0:
            for g in grads.values():
                g.mul_(coef)
    h.step += 1
    for name, p in h.trainable().items():
        _adam_inplace(p, grads[name], h.m[name], h.v[name], lr, h.step)
    return total


def hyper_round_torch(h: FusedHyperNet, clients: Sequence[int], client_rows: torch.Tensor, lr: float,
                      clip: float, mode: str = "sequential") -> List[float]:
    """``client_rows[j]`` is the update of ``clients[j]``.  ``sequential`` = reference semantics;
    ``batched`` = one mini-batch step over all clients (all-gather -> GEMM form, a different
    algorithm: mean of the per-client gradients, one Adam step)."""
    norms: List[float] = []
    if mode == "sequential":
        for j, c in enumerate(clients):
            norms.append(hyper_step_torch(h, int(c), client_rows[j], lr, clip))
        return norms
    acc: Optional[Dict[str, torch.Tensor]] = None
    for j, c in enumerate(clients):
        g = hyper_grads_torch(h, int(c), client_rows[j])
        g.pop("_delta_sq")
        if acc is None:
            acc = {k: v.clone() for k, v in g.items()}
        else:
            for k in acc:
                acc[k].add_(g[k])
    assert acc is not None
    n = float(len(clients))
    for k in acc:
        acc[k].div_(n)
    total = math.sqrt(sum(float((g.double() ** 2).sum()) for g in acc.values()))
    if clip and clip > 0:
        coef = min(1.0, clip / (total + 1e-6))
        if coef < 1.0:
            for g in acc.values():
                g.mul_(coef)
    h.step += 1
    for name, p in h.trainable().items():
        _adam_inplace(p, acc[name], h.m[name], h.v[name], lr, h.step)
    return [total]
