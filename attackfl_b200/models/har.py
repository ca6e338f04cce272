"""HAR model: ``PositionalEncoding`` + ``TransformerClassifier`` (T=561, d=64, 4 heads, 2 layers).

Reference parity: ``/root/reference/src/Model.py:420-433`` (sinusoidal table ``pe [1,600,d]``
registered as a buffer, so it is part of the ``state_dict``) and ``:435-458`` (Conv1d(1->64,k3)
over the 561 features, +PE, two seq-first ``nn.TransformerEncoderLayer(d, heads, ff=256, p=0.1)``,
mean over T, Linear-ReLU-Dropout(0.3)-Linear -> 6 classes).  The reference's *client* HAR path is
broken (``src/RpcClient.py:157-164`` falls into ``raise``); here it works.
"""
from __future__ import annotations


import torch
import torch.nn as nn

HAR_FEATURES = 561
HAR_CLASSES = 6


class PositionalEncoding(nn.Module):
    def __init__(self, d_model: int, max_len: int = 600):
        super().__init__()
        pos = torch.arange(max_len, dtype=torch.float32)[:, None]
        freq = torch.exp(torch.arange(0, d_model, 2, dtype=torch.float32)
                         * (-torch.log(torch.tensor(10000.0)) / d_model))
        table = torch.zeros(max_len, d_model)
        table[:, 0::2] = torch.sin(pos * freq)
        table[:, 1::2] = torch.cos(pos * freq)
        self.register_buffer("pe", table[None])        # [1, max_len, d_model]

    def forward(self, x: torch.Tensor) -> torch.Tensor:   # x: [B, T, d]
        return x + self.pe[:, :x.size(1), :]


class TransformerClassifier(nn.Module):
    def __init__(self, d_model: int = 64, nhead: int = 4, num_layers: int = 2, num_classes: int = HAR_CLASSES):
        super().__init__()
        self.conv = nn.Conv1d(1, d_model, kernel_size=3, padding=1)
        self.pe = PositionalEncoding(d_model)
        layer = nn.TransformerEncoderLayer(d_model=d_model, nhead=nhead, dim_feedforward=256, dropout=0.1)
        self.transformer = nn.TransformerEncoder(layer, num_layers=num_layers, enable_nested_tensor=False)
        self.classifier = nn.Sequential(nn.Linear(d_model, 64), nn.ReLU(), nn.Dropout(0.3),
                                        nn.Linear(64, num_classes))

    def forward(self, x: torch.Tensor) -> torch.Tensor:   # x: [B, 1, 561]
        h = self.conv(x).transpose(1, 2)                  # [B, T, d]
        h = self.pe(h).transpose(0, 1)                    # [T, B, d] (seq-first encoder)
        h = self.transformer(h).mean(dim=0)
        return self.classifier(h)
