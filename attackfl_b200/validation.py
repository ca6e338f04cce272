"""Server-side validation: ICU ROC-AUC (plain and hyper-pooled), HAR accuracy; NaN => round fails.

Reference parity: ``/root/reference/src/Validation.py:50-67`` (``test``), ``:92-122`` (``test_icu``:
no-grad forward over the test set in batches of 64, NaN in outputs => ``False``, sklearn
``roc_curve``+``auc``, prints and logs ``ROC_AUC: x.xxxx``), ``:124-136`` (``test_har``),
``:138-145,178-213`` (``test_hyper*``: one full test pass per client with weights emitted by the
hyper-network, all clients' outputs pooled into a single ROC-AUC).

AUC is computed on the device by sort + tie-aware trapezoid (equal to sklearn's ``auc(roc_curve)``
up to fp rounding) so the scores never travel to the host; only the scalar does.
"""
from __future__ import annotations

from typing import Optional

import torch

from .log import print_with_color


def roc_auc(scores: torch.Tensor, labels: torch.Tensor) -> float:
    """Tie-aware ROC-AUC == ``sklearn.metrics.auc(*roc_curve(labels, scores)[:2])``."""
    s = scores.reshape(-1).to(torch.float64)
    y = labels.reshape(-1).to(torch.float64)
    n_pos = y.sum()
    n_neg = y.numel() - n_pos
    if n_pos == 0 or n_neg == 0:
        return float("nan")
    order = torch.argsort(s, descending=True, stable=True)
    s, y = s[order], y[order]
    tp = torch.cumsum(y, 0)
    fp = torch.cumsum(1 - y, 0)
    last = torch.ones_like(s, dtype=torch.bool)
    last[:-1] = s[1:] != s[:-1]                       # last element of every tie group
    tp, fp = tp[last], fp[last]
    tp = torch.cat([tp.new_zeros(1), tp])
    fp = torch.cat([fp.new_zeros(1), fp])
    area = torch.trapezoid(tp, fp)
    return float(area / (n_pos * n_neg))


class Validator:
    """``forward_fn(row_or_None, vitals, labs) -> probs`` abstracts torch modules vs the native kernel."""

    def __init__(self, data_name: str, logger=None, quiet: bool = False, native: bool = False):
        self.data_name = data_name
        self.native = native             # ROC-AUC through the native library (``ops.flat.roc_auc``) when it fits
        self.logger = logger
        self.quiet = quiet
        self.last_metric: Optional[float] = None

    def _emit(self, text: str) -> None:
        if self.quiet:
            return
        print(text, flush=True)
        if self.logger is not None:
            self.logger.log_info(text)

    def score_icu(self, outputs: torch.Tensor, labels: torch.Tensor) -> bool:
        if self.native and outputs.numel() <= 8192:
            # one launch (sort + tie-aware trapezoid in a single block) and one 16-byte read instead of ~20 torch
            # launches and three host syncs
            from .ops import flat
            auc, n_nan = flat.roc_auc(outputs, labels)[:2].tolist()
            if n_nan:
                print_with_color("NaN detected in output, training false", "yellow")
                return False
            self.last_metric = auc
            self._emit(f"ROC_AUC: {auc:.4f}")
            return True
        if torch.isnan(outputs).any():
            print_with_color("NaN detected in output, training false", "yellow")
            return False
        auc = roc_auc(outputs, labels)
        self.last_metric = auc
        self._emit(f"ROC_AUC: {auc:.4f}")
        return True

    def score_har(self, logits: torch.Tensor, labels: torch.Tensor) -> bool:
        if torch.isnan(logits).any():
            print_with_color("NaN detected in output, training false", "yellow")
            return False
        acc = (logits.argmax(dim=1) == labels).float().mean().item()
        self.last_metric = acc
        self._emit(f"Test Accuracy: {acc:.4f}")
        return True
