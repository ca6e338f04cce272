"""Datasets: reference-compatible ``ICUData``, gzip-pickle loaders, synthetic generators, device packing.

Reference parity: ``/root/reference/src/Model.py:9-24`` (``ICUData(dataframe, vitals_cols, labs_cols,
label_col)`` -> fp32 tensors, ``__getitem__ -> (vitals[7], labs[16], label)``), loaders at
``src/RpcClient.py:155-169`` (``train_dataset.pkl.gz`` in the CWD, ``data/icu_har_train_ds.pkl.gz``)
and ``src/Validation.py:32-37`` (``data/test_dataset.pkl.gz``, ``data/icu_har_test_ds.pkl.gz``).
The dataset blobs are absent from the reference mount (``.MISSING_LARGE_BLOBS``), so synthetic
generators of the same shapes are provided; real blobs are picked up when present.

Device packing: ICU rows are stored as one ``[N, 24]`` fp32 matrix (vitals 0..6, labs 7..22,
label 23) so the training kernel gathers one 96-byte record per sample.
"""
from __future__ import annotations

import gzip
import io
import os
import pickle
import random
from typing import Optional

import torch
from torch.utils.data import Dataset

from .models.icu import LABS_DIM, VITALS_DIM, MASK_VALUE
from .models.har import HAR_CLASSES, HAR_FEATURES

ICU_RECORD = VITALS_DIM + LABS_DIM + 1   # 24 floats


class ICUData(Dataset):
    """Drop-in for the reference's ``ICUData``; also constructible from tensors."""

    def __init__(self, dataframe=None, vitals_cols=None, labs_cols=None, label_col=None, *,
                 vitals: Optional[torch.Tensor] = None, labs: Optional[torch.Tensor] = None,
                 labels: Optional[torch.Tensor] = None):
        if dataframe is not None:
            vitals = torch.tensor(dataframe[vitals_cols].values, dtype=torch.float32)
            labs = torch.tensor(dataframe[labs_cols].values, dtype=torch.float32)
            labels = torch.tensor(dataframe[label_col].values, dtype=torch.float32)
        self.vitals = vitals.to(torch.float32)
        self.labs = labs.to(torch.float32)
        self.labels = labels.to(torch.float32)

    def __len__(self):
        return len(self.labels)

    def __getitem__(self, idx):
        return self.vitals[idx], self.labs[idx], self.labels[idx]

    def packed(self) -> torch.Tensor:
        return torch.cat([self.vitals, self.labs, self.labels[:, None]], dim=1).contiguous()


class HARData(Dataset):
    """``(x [1,561] float, y long)`` samples (shape inferred from Model.py:438,451 / client.py:122-126)."""

    def __init__(self, x: torch.Tensor, y: torch.Tensor):
        self.x = x.to(torch.float32)
        self.y = y.to(torch.long)

    def __len__(self):
        return len(self.y)

    def __getitem__(self, idx):
        return self.x[idx], self.y[idx]


class _CompatUnpickler(pickle.Unpickler):
    """Resolves ``ICUData`` pickled from ``__main__`` / ``src.Model`` / ``src.Utils`` to our class."""

    def find_class(self, module, name):
        if name == "ICUData":
            return ICUData
        if name == "HARData":
            return HARData
        return super().find_class(module, name)


def load_pkl_gz(path: str):
    with gzip.open(path, "rb") as fh:
        return _CompatUnpickler(io.BytesIO(fh.read())).load()


def save_pkl_gz(obj, path: str) -> None:
    os.makedirs(os.path.dirname(os.path.abspath(path)) or ".", exist_ok=True)
    with gzip.open(path, "wb") as fh:
        pickle.dump(obj, fh)


# ---------------------------------------------------------------------------- synthetic data
def synthetic_icu(n: int, seed: int = 0, mask_frac: float = 0.03) -> ICUData:
    """ICU-shaped data with a learnable, non-linear signal (AUC of a good model ~0.85-0.9)."""
    g = torch.Generator().manual_seed(1000003 * seed + 17)
    x = torch.randn(n, VITALS_DIM + LABS_DIM, generator=g)
    wg = torch.Generator().manual_seed(424243)           # the concept is shared by train and test
    w1 = torch.randn(VITALS_DIM + LABS_DIM, generator=wg) / 3.0
    w2 = torch.randn(VITALS_DIM + LABS_DIM, 4, generator=wg) / 3.0
    logit = x @ w1 + torch.tanh(x @ w2).sum(1) * 0.8 - 0.9
    logit = logit + 0.5 * torch.randn(n, generator=g)
    labels = (torch.rand(n, generator=g) < torch.sigmoid(1.5 * logit)).to(torch.float32)
    if mask_frac > 0:
        m = torch.rand(x.shape, generator=g) < mask_frac
        x = torch.where(m, torch.full_like(x, MASK_VALUE), x)
    return ICUData(vitals=x[:, :VITALS_DIM].contiguous(), labs=x[:, VITALS_DIM:].contiguous(), labels=labels)


def synthetic_har(n: int, seed: int = 0) -> HARData:
    g = torch.Generator().manual_seed(7919 * seed + 3)
    cg = torch.Generator().manual_seed(99991)
    proto = torch.randn(HAR_CLASSES, HAR_FEATURES, generator=cg)
    y = torch.randint(0, HAR_CLASSES, (n,), generator=g)
    x = 0.6 * proto[y] + torch.randn(n, HAR_FEATURES, generator=g)
    return HARData(x[:, None, :].contiguous(), y)


def load_train(data_name: str, rows: int, seed: int = 0):
    """Real blob if present (same paths as the reference), else synthetic of ``rows`` samples."""
    if data_name == "ICU":
        for p in ("train_dataset.pkl.gz", "data/train_dataset.pkl.gz"):
            if os.path.exists(p):
                return load_pkl_gz(p), p
        return synthetic_icu(rows, seed), "synthetic"
    if data_name == "HAR":
        p = "data/icu_har_train_ds.pkl.gz"
        if os.path.exists(p):
            return load_pkl_gz(p), p
        return synthetic_har(rows, seed), "synthetic"
    raise ValueError(f"Data name '{data_name}' is not valid.")


def load_test(data_name: str, rows: int, seed: int = 1):
    if data_name == "ICU":
        p = "data/test_dataset.pkl.gz"
        if os.path.exists(p):
            return load_pkl_gz(p), p
        return synthetic_icu(rows, seed + 101), "synthetic"
    if data_name == "HAR":
        p = "data/icu_har_test_ds.pkl.gz"
        if os.path.exists(p):
            return load_pkl_gz(p), p
        return synthetic_har(rows, seed + 101), "synthetic"
    raise ValueError(f"Data name '{data_name}' is not valid.")


def as_har(ds) -> HARData:
    if isinstance(ds, HARData):
        return ds
    xs, ys = zip(*[ds[i] for i in range(len(ds))])
    return HARData(torch.stack([torch.as_tensor(x) for x in xs]), torch.as_tensor(ys))


# ---------------------------------------------------------------------------- per-round sampling
def draw_num_data(rng: random.Random, lo: int, hi: int) -> int:
    """``random.randrange(lo, hi + 1)`` — src/RpcClient.py:97."""
    return rng.randrange(lo, hi + 1)


def make_round_indices(n_rows: int, sizes, epochs: int, generator: torch.Generator,
                       device="cpu") -> torch.Tensor:
    """Per client: a random subset of ``size`` rows (``random.sample``, RpcClient.py:166) and an
    independent shuffle per epoch (``DataLoader(shuffle=True)``, :169).

    Returns int32 ``[C, epochs, max_size]`` (entries past ``size`` are -1).  Generated with
    ``argsort(rand)`` on ``device`` so nothing crosses PCIe per round.
    """
    c = len(sizes)
    mx = int(max(sizes))
    dev = torch.device(device)
    sizes_t = torch.as_tensor(sizes, device=dev, dtype=torch.long)
    subset = torch.rand(c, n_rows, generator=generator, device=dev).argsort(dim=1)[:, :mx]   # [C, mx]
    col = torch.arange(mx, device=dev)[None, :]
    valid = col < sizes_t[:, None]
    keys = torch.rand(c, epochs, mx, generator=generator, device=dev)
    keys = torch.where(valid[:, None, :], keys, torch.full_like(keys, 2.0))   # invalid sort last
    order = keys.argsort(dim=2)
    idx = torch.gather(subset[:, None, :].expand(c, epochs, mx), 2, order)
    idx = torch.where(valid[:, None, :].expand_as(idx), idx, torch.full_like(idx, -1))
    return idx.to(torch.int32).contiguous()
