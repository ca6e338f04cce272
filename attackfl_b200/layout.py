"""FlatLayout: ``state_dict`` keys/shapes <-> offsets in one flat fp32 vector, plus ``.pth`` I/O.

Reference parity: the reference moves whole ``state_dict``s around (pickled, ``server.py:614``,
``src/RpcClient.py:111-114``) and flattens them ad hoc in ``state_dict`` key order
(``src/Utils.py:225-226,250-255,360-361``).  Here the flat vector *is* the storage: every client
model is one row of a ``[C, D_pad]`` fp32 arena and the per-key tensors are views into the row.

Layout rules: tensors appear in ``state_dict`` order, each tensor starts on a
128-byte boundary (32 floats) and ``D_pad`` is a multiple of 1024 floats so that slices owned by up
to 8 GPUs, ``multimem`` v4 accesses and TMA boxes stay aligned.  Padding is zero in every row
(neutral for sums, Gram matrices and medians) and is stripped on export, so ``{model}.pth`` files
are byte-compatible with ``torch.load`` + ``load_state_dict`` of the reference's modules
(``server.py:549-553``).
"""
from __future__ import annotations

from collections import OrderedDict
from dataclasses import dataclass
from typing import Dict, Iterable, List, Mapping, Optional, Sequence, Tuple

import torch

TENSOR_ALIGN = 32      # floats (128 B)
ROW_ALIGN = 1024       # floats


def _round_up(x: int, m: int) -> int:
    return (x + m - 1) // m * m


@dataclass(frozen=True)
class Entry:
    name: str
    shape: Tuple[int, ...]
    numel: int
    offset: int          # in floats, inside the padded row
    dense_offset: int    # offset in the unpadded concatenation (reference's state_dict_to_vector)


class FlatLayout:
    def __init__(self, named_shapes: Iterable[Tuple[str, Sequence[int]]]):
        entries: List[Entry] = []
        off = 0
        dense = 0
        for name, shape in named_shapes:
            shape = tuple(int(s) for s in shape)
            n = 1
            for s in shape:
                n *= s
            off = _round_up(off, TENSOR_ALIGN)
            entries.append(Entry(name, shape, n, off, dense))
            off += n
            dense += n
        self.entries: List[Entry] = entries
        self.by_name: Dict[str, Entry] = {e.name: e for e in entries}
        self.numel: int = dense                       # D (reference's flat length)
        self.padded: int = _round_up(max(off, 1), ROW_ALIGN)   # D_pad
        self._dense_index: Optional[torch.Tensor] = None

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_state_dict(cls, sd: Mapping[str, torch.Tensor]) -> "FlatLayout":
        for k, v in sd.items():
            if not torch.is_floating_point(v):
                raise TypeError(f"FlatLayout: non-float entry '{k}' ({v.dtype}) is not supported")
        return cls((k, v.shape) for k, v in sd.items())

    @classmethod
    def from_module(cls, module: torch.nn.Module) -> "FlatLayout":
        return cls.from_state_dict(module.state_dict())

    # ------------------------------------------------------------------ properties
    def keys(self) -> List[str]:
        return [e.name for e in self.entries]

    def offset(self, name: str) -> int:
        return self.by_name[name].offset

    def __len__(self) -> int:
        return len(self.entries)

    def dense_index(self, device=None) -> torch.Tensor:
        """int64 index of length D mapping dense position -> padded position."""
        if self._dense_index is None:
            parts = [torch.arange(e.offset, e.offset + e.numel, dtype=torch.long) for e in self.entries]
            self._dense_index = torch.cat(parts) if parts else torch.zeros(0, dtype=torch.long)
        idx = self._dense_index
        return idx.to(device) if device is not None else idx

    def segments(self, device=None) -> torch.Tensor:
        """``[T, 2]`` int32 (offset, numel) per tensor — used by the ``layer-fro`` metric."""
        t = torch.tensor([[e.offset, e.numel] for e in self.entries], dtype=torch.int32)
        return t.to(device) if device is not None else t

    def valid_mask(self, device=None) -> torch.Tensor:
        m = torch.zeros(self.padded, dtype=torch.bool)
        m[self.dense_index()] = True
        return m.to(device) if device is not None else m

    # ------------------------------------------------------------------ conversions
    def flatten(self, sd: Mapping[str, torch.Tensor], out: Optional[torch.Tensor] = None,
                device=None, strict_shapes: bool = False) -> torch.Tensor:
        """state_dict -> padded fp32 row (zeros in the padding).  ``strict_shapes``: refuse a tensor whose element count
        matches but whose shape does not (used for files that come from outside, ``load_pth``)."""
        if out is None:
            out = torch.zeros(self.padded, dtype=torch.float32, device=device)
        else:
            out.zero_()
        for e in self.entries:
            src = sd[e.name]
            if tuple(src.shape) != e.shape and (strict_shapes or src.numel() != e.numel):
                raise ValueError(f"shape mismatch for '{e.name}': {tuple(src.shape)} vs {e.shape}")
            out[e.offset:e.offset + e.numel].copy_(src.detach().reshape(-1).to(torch.float32))
        return out

    def views(self, row: torch.Tensor) -> "OrderedDict[str, torch.Tensor]":
        """padded row -> OrderedDict of *views* shaped like the state_dict (no copy)."""
        assert row.dim() == 1 and row.numel() >= self.padded
        return OrderedDict((e.name, row[e.offset:e.offset + e.numel].view(e.shape)) for e in self.entries)

    def batched_views(self, rows: torch.Tensor) -> "OrderedDict[str, torch.Tensor]":
        """``[C, D_pad]`` -> name -> ``[C, *shape]`` views (for multi-client ``torch.func`` / bmm)."""
        assert rows.dim() == 2
        c = rows.shape[0]
        return OrderedDict(
            (e.name, rows[:, e.offset:e.offset + e.numel].view((c,) + e.shape)) for e in self.entries)

    def unflatten(self, row: torch.Tensor, clone: bool = True, device=None) -> "OrderedDict[str, torch.Tensor]":
        sd = self.views(row)
        if clone or device is not None:
            sd = OrderedDict((k, (v.to(device) if device is not None else v).clone()) for k, v in sd.items())
        return sd

    def to_dense(self, rows: torch.Tensor) -> torch.Tensor:
        """padded row(s) -> reference-order dense vector(s) of length D."""
        return rows.index_select(-1, self.dense_index(rows.device))

    def from_dense(self, dense: torch.Tensor) -> torch.Tensor:
        shape = dense.shape[:-1] + (self.padded,)
        out = torch.zeros(shape, dtype=torch.float32, device=dense.device)
        out.index_copy_(-1, self.dense_index(dense.device), dense.to(torch.float32))
        return out

    # ------------------------------------------------------------------ .pth I/O
    def save_pth(self, row: torch.Tensor, path: str) -> None:
        """Writes ``{model}.pth`` exactly like ``torch.save(final_state_dict, ...)`` (server.py:553):
        a plain dict of CPU fp32 tensors keyed by the module's state_dict names."""
        sd = self.unflatten(row.detach(), clone=True, device="cpu")
        tmp = path + ".tmp"
        torch.save(dict(sd), tmp)
        import os
        os.replace(tmp, path)

    def load_pth(self, path: str, device=None) -> torch.Tensor:
        sd = torch.load(path, map_location="cpu", weights_only=True)
        missing = [k for k in self.keys() if k not in sd]
        if missing:
            raise KeyError(f"{path}: missing keys {missing[:4]}{'...' if len(missing) > 4 else ''}")
        return self.flatten(sd, device=device, strict_shapes=True)

    def describe(self) -> str:
        return f"FlatLayout(T={len(self.entries)}, D={self.numel}, D_pad={self.padded})"
