"""Per-phase timers: CUDA events on the launching stream (device time), wall clock on CPU.

The reference has no timers at all.  ``PhaseTimer`` records train / craft / aggregate /
hyper / validate phases per round; multi-GPU reports take the max over ranks (``parallel.ctx``)."""
from __future__ import annotations

import time
from collections import OrderedDict
from typing import Dict, List, Tuple

import torch


class PhaseTimer:
    def __init__(self, device, nvtx: bool = True):
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        self.nvtx = bool(nvtx) and self.cuda
        self._open: List[Tuple[str, object]] = []
        self._pending: List[Tuple[str, object, object]] = []
        self.totals: "OrderedDict[str, float]" = OrderedDict()

    def start(self, name: str) -> None:
        if self.cuda:
            if self.nvtx:
                torch.cuda.nvtx.range_push(f"afl:{name}")
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
            if self.nvtx:
                torch.cuda.nvtx.range_pop()
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


class Watchdog:
    """Rank-health watchdog (SURVEY 5.3 plan): if a round does not finish within ``timeout_s`` the rank logs the
    phase it is stuck in and exits with code 75, which takes the whole torchrun job down instead of leaving the
    other ranks spinning in a collective.  The in-kernel cross-GPU flag waits carry their own ``clock64`` bound."""

    def __init__(self, timeout_s: float, on_expire=None):
        import threading
        self.timeout_s, self.on_expire = float(timeout_s), on_expire
        self._timer = None
        self._threading = threading
        self.expired = False

    def arm(self, what: str) -> None:
        self.disarm()
        if self.timeout_s <= 0:
            return
        self._timer = self._threading.Timer(self.timeout_s, self._fire, args=(what,))
        self._timer.daemon = True
        self._timer.start()

    def disarm(self) -> None:
        if self._timer is not None:
            self._timer.cancel()
            self._timer = None

    def _fire(self, what: str) -> None:
        self.expired = True
        if self.on_expire is not None:
            self.on_expire(what)
            return
        import os
        import sys
        print(f"[watchdog] {what} exceeded {self.timeout_s:.1f} s; aborting this rank", file=sys.stderr, flush=True)
        os._exit(75)
