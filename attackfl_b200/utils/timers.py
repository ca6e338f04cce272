"""Per-phase timers: CUDA events on the launching stream (device time), wall clock on CPU.

The reference has no timers at all.  ``PhaseTimer`` records train / craft / aggregate /
hyper / validate phases per round; multi-GPU reports take the max over ranks (``parallel.ctx``)."""
from __future__ import annotations

import time
from collections import OrderedDict
from typing import Dict, List, Tuple

import torch


class PhaseTimer:
    def __init__(self, device):
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self._open: List[Tuple[str, object]] = []
        self._pending: List[Tuple[str, object, object]] = []
        self.totals: "OrderedDict[str, float]" = OrderedDict()

    def start(self, name: str) -> None:
        if self.cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream(self.device))
            self._open.append((name, ev))
        else:
            self._open.append((name, time.perf_counter()))

    def stop(self) -> None:
        name, t0 = self._open.pop()
        if self.cuda:
            ev = torch.cuda.Event(enable_timing=True)
            ev.record(torch.cuda.current_stream(self.device))
            self._pending.append((name, t0, ev))
        else:
            self.totals[name] = self.totals.get(name, 0.0) + (time.perf_counter() - t0) * 1e3

    class _Ctx:
        def __init__(self, timer, name):
            self.t, self.n = timer, name

        def __enter__(self):
            self.t.start(self.n)

        def __exit__(self, *a):
            self.t.stop()

    def phase(self, name: str) -> "_Ctx":
        return PhaseTimer._Ctx(self, name)

    def collect(self) -> Dict[str, float]:
        """Milliseconds per phase since the last ``collect`` (synchronises the recorded events only)."""
        if self.cuda:
            for name, e0, e1 in self._pending:
                e1.synchronize()
                self.totals[name] = self.totals.get(name, 0.0) + e0.elapsed_time(e1)
            self._pending = []
        out = dict(self.totals)
        self.totals = OrderedDict()
        return out
