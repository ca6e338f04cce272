"""ICU models: ``CNNModel``, ``RNNModel``, ``TransformerBlock``, ``TransformerModel``.

Reference parity: ``/root/reference/src/Model.py:27-88`` (CNN), ``:91-163`` (RNN), ``:166-191``
(block), ``:194-246`` (Transformer).  The modules are *key- and shape-compatible* with the
reference (same ``state_dict`` names, order and default initialisers), so ``{model}.pth`` files
load in either code base.  They are the single-client oracle / general-T implementation; the hot
path is the multi-client native trainer (``attackfl_b200.ops.train``), which executes the T=1
closed forms documented in SURVEY §2.7 directly on rows of the flat arena.

Each model also exposes ``BRANCHES`` / ``spec()`` metadata that the op-program builder
(``models/programs.py``) turns into the instruction list of the persistent training kernel.
"""
from __future__ import annotations

from typing import Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

VITALS_DIM = 7
LABS_DIM = 16
BRANCHES: Tuple[Tuple[str, int], ...] = (("vitals", VITALS_DIM), ("labs", LABS_DIM))
MASK_VALUE = -2.0


class CNNModel(nn.Module):
    """Two Conv1d(k=3,p=1) towers 1->32->64->128 + AdaptiveAvgPool1d(4) + MLP 1024-128-64-32-1."""

    CHANNELS = (1, 32, 64, 128)
    POOL = 4
    DROPOUT = 0.3

    def __init__(self):
        super().__init__()
        for branch, _ in BRANCHES:
            for i in range(3):
                setattr(self, f"{branch}_conv{i + 1}",
                        nn.Conv1d(self.CHANNELS[i], self.CHANNELS[i + 1], kernel_size=3, padding=1))
            setattr(self, f"{branch}_pool", nn.AdaptiveAvgPool1d(self.POOL))
            setattr(self, f"{branch}_dropout", nn.Dropout(self.DROPOUT))
        width = self.CHANNELS[-1] * self.POOL * len(BRANCHES)
        self.fc1 = nn.Linear(width, 128)
        self.fc2 = nn.Linear(128, 64)
        self.fc3 = nn.Linear(64, 32)
        self.output = nn.Linear(32, 1)

    def _tower(self, branch: str, x: torch.Tensor) -> torch.Tensor:
        x = x.unsqueeze(1)
        for i in (1, 2, 3):
            x = F.relu(getattr(self, f"{branch}_conv{i}")(x))
        x = getattr(self, f"{branch}_pool")(x).flatten(1)
        return getattr(self, f"{branch}_dropout")(x)

    def forward(self, vitals: torch.Tensor, labs: torch.Tensor) -> torch.Tensor:
        x = torch.cat([self._tower("vitals", vitals), self._tower("labs", labs)], dim=1)
        for fc in (self.fc1, self.fc2, self.fc3):
            x = F.relu(fc(x))
        return torch.sigmoid(self.output(x))


class RNNModel(nn.Module):
    """Two 3-layer bidirectional GRU stacks (H=32) + LayerNorm(64) + MLP 128-32-16-1.

    The reference always feeds 2-D inputs, i.e. sequence length 1 (Model.py:125-126); 3-D inputs
    ``[B, T, F]`` are accepted as well and use the last time step, like the reference.
    """

    def __init__(self, vitals_input_dim: int = VITALS_DIM, labs_input_dim: int = LABS_DIM,
                 hidden_dim: int = 32, dropout_rate: float = 0.3):
        super().__init__()
        self.mask_value = MASK_VALUE
        self.hidden_dim = hidden_dim
        for branch, in_dim in (("vitals", vitals_input_dim), ("labs", labs_input_dim)):
            dims = (in_dim, 2 * hidden_dim, 2 * hidden_dim)
            for i, d in enumerate(dims):
                setattr(self, f"{branch}_gru{i + 1}",
                        nn.GRU(d, hidden_dim, batch_first=True, bidirectional=True))
            setattr(self, f"{branch}_ln", nn.LayerNorm(2 * hidden_dim))
            setattr(self, f"{branch}_dropout", nn.Dropout(dropout_rate))
        self.fc1 = nn.Linear(4 * hidden_dim, hidden_dim)
        self.fc2 = nn.Linear(hidden_dim, hidden_dim // 2)
        self.output = nn.Linear(hidden_dim // 2, 1)

    def _stack(self, branch: str, x: torch.Tensor) -> torch.Tensor:
        x = torch.where(x == self.mask_value, torch.zeros_like(x), x)
        if x.dim() == 2:
            x = x.unsqueeze(1)
        for i in (1, 2, 3):
            x, _ = getattr(self, f"{branch}_gru{i}")(x)
        x = getattr(self, f"{branch}_ln")(x[:, -1, :])
        return getattr(self, f"{branch}_dropout")(x)

    def forward(self, vitals: torch.Tensor, labs: torch.Tensor) -> torch.Tensor:
        x = torch.cat([self._stack("vitals", vitals), self._stack("labs", labs)], dim=1)
        x = F.relu(self.fc1(x))
        x = F.relu(self.fc2(x))
        return torch.sigmoid(self.output(x))


class TransformerBlock(nn.Module):
    """Post-LN block: MHA -> add&norm -> FFN(GELU) -> add&norm  (Model.py:166-191)."""

    def __init__(self, input_dim: int, num_heads: int, ff_dim: int, dropout_rate: float = 0.1):
        super().__init__()
        self.attention = nn.MultiheadAttention(embed_dim=input_dim, num_heads=num_heads,
                                               dropout=dropout_rate, batch_first=True)
        self.attention_norm = nn.LayerNorm(input_dim)
        self.dropout1 = nn.Dropout(dropout_rate)
        self.ffn = nn.Sequential(nn.Linear(input_dim, ff_dim), nn.GELU(), nn.Dropout(dropout_rate),
                                 nn.Linear(ff_dim, input_dim))
        self.ffn_norm = nn.LayerNorm(input_dim)
        self.dropout2 = nn.Dropout(dropout_rate)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        a, _ = self.attention(x, x, x, need_weights=False)
        x = self.attention_norm(x + self.dropout1(a))
        return self.ffn_norm(x + self.dropout2(self.ffn(x)))


class TransformerModel(nn.Module):
    """Linear->GELU->TransformerBlock(64, heads, ff)->LayerNorm per branch; MLP 128-64-32-1.

    ``d_model`` / ``depth`` are exposed (the reference hard-codes 64 / 1, Model.py:202,211) so the
    ~100 M-parameter aggregation sweep of BASELINE config 5 can instantiate a scaled variant; the
    defaults reproduce the reference exactly (47,693 parameters, 38 tensors).
    """

    def __init__(self, vitals_input_dim: int = VITALS_DIM, labs_input_dim: int = LABS_DIM,
                 num_heads: int = 4, ff_dim: int = 6, d_model: int = 64):
        super().__init__()
        self.d_model = d_model
        for branch, in_dim in (("vitals", vitals_input_dim), ("labs", labs_input_dim)):
            setattr(self, f"{branch}_dense", nn.Linear(in_dim, d_model))
            setattr(self, f"{branch}_transformer", TransformerBlock(d_model, num_heads, ff_dim))
            # parameter-free, unused in forward (kept so the module tree matches Model.py:204-207)
            setattr(self, f"{branch}_pool", nn.Sequential(nn.AdaptiveAvgPool1d(1), nn.AdaptiveMaxPool1d(1)))
            setattr(self, f"{branch}_bn", nn.LayerNorm(d_model))
        self.fc1 = nn.Linear(2 * d_model, d_model)
        self.dropout = nn.Dropout(0.3)
        self.fc2 = nn.Linear(d_model, d_model // 2)
        self.output = nn.Linear(d_model // 2, 1)

    def _branch(self, branch: str, x: torch.Tensor) -> torch.Tensor:
        x = F.gelu(getattr(self, f"{branch}_dense")(x)).unsqueeze(1)
        x = getattr(self, f"{branch}_transformer")(x).squeeze(1)
        return getattr(self, f"{branch}_bn")(x)

    def forward(self, vitals: torch.Tensor, labs: torch.Tensor) -> torch.Tensor:
        x = torch.cat([self._branch("vitals", vitals), self._branch("labs", labs)], dim=1)
        x = self.dropout(F.gelu(self.fc1(x)))
        x = F.gelu(self.fc2(x))
        return torch.sigmoid(self.output(x))
