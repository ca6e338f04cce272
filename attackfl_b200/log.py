"""Coloured stdout, ``app.log`` with the reference's line format, and ``metrics.jsonl``.

Reference parity: ``/root/reference/src/Log.py:1-44`` (``Colors``, ``Logger``, ``print_with_color``).
``app.log`` lines keep the exact format ``%(asctime)s - my_logger - LEVEL - msg`` so log scrapers
written against the reference keep working.  ``MetricsWriter`` is new: one JSON
object per round (phase timings max-over-ranks, ROC-AUC, gamma, attackers, NaN flags).
"""
from __future__ import annotations

import json
import logging
import os
import time
from typing import Any, Dict, Optional


class Colors:
    COLORS = {
        "header": "\033[95m",
        "blue": "\033[94m",
        "green": "\033[92m",
        "yellow": "\033[93m",
        "red": "\033[91m",
        "end": "\033[0m",
    }


_QUIET = bool(int(os.environ.get("ATTACKFL_QUIET", "0")))


def set_quiet(flag: bool) -> None:
    global _QUIET
    _QUIET = bool(flag)


def print_with_color(text: str, color: str) -> None:
    if _QUIET:
        return
    code = Colors.COLORS.get(str(color).lower(), Colors.COLORS["end"])
    print(f"{code}{text}{Colors.COLORS['end']}", flush=True)


class Logger:
    """File logger; one handler per path (the reference leaks a handler per instance)."""

    def __init__(self, log_path: str):
        self.logger = logging.getLogger("my_logger")
        self.logger.setLevel(logging.DEBUG)
        self.logger.propagate = False
        apath = os.path.abspath(log_path)
        os.makedirs(os.path.dirname(apath) or ".", exist_ok=True)
        for h in list(self.logger.handlers):
            if isinstance(h, logging.FileHandler) and h.baseFilename == apath:
                break
        else:
            fh = logging.FileHandler(apath)
            fh.setLevel(logging.DEBUG)
            fh.setFormatter(logging.Formatter("%(asctime)s - %(name)s - %(levelname)s - %(message)s"))
            self.logger.addHandler(fh)

    def log_info(self, message: str) -> None:
        self.logger.info(message)

    def log_warning(self, message: str) -> None:
        self.logger.warning(message)

    def log_error(self, message: str) -> None:
        self.logger.error(message)

    def close(self) -> None:
        for h in list(self.logger.handlers):
            h.close()
            self.logger.removeHandler(h)


class NullLogger:
    def log_info(self, message: str) -> None: ...
    def log_warning(self, message: str) -> None: ...
    def log_error(self, message: str) -> None: ...
    def close(self) -> None: ...


class MetricsWriter:
    """Append-only JSONL of per-round metrics (new; the reference persists only ROC_AUC lines)."""

    def __init__(self, path: Optional[str]):
        self.path = path
        self._fh = open(path, "a") if path else None

    def write(self, record: Dict[str, Any]) -> None:
        if self._fh is None:
            return
        rec = dict(record)
        rec.setdefault("ts", time.time())
        self._fh.write(json.dumps(rec, default=float) + "\n")
        self._fh.flush()

    def close(self) -> None:
        if self._fh is not None:
            self._fh.close()
            self._fh = None
