"""Tracing / profiling helpers (the reference has no profiler hooks, timers
or per-job timing at all).  `StageTimer` brackets pipeline stages with CUDA events (device time, no
host sync until read) and NVTX ranges (`torch.cuda.nvtx`), and `Watchdog` turns a wedged device
call into a Python exception instead of a hung worker (section 5.3)."""
from __future__ import annotations

import contextlib
import threading
import time
from typing import Dict, List, Tuple

import torch


class StageTimer:
    def __init__(self, enabled: bool = True):
        self.enabled = enabled and torch.cuda.is_available()
        self._events: List[Tuple[str, "torch.cuda.Event", "torch.cuda.Event"]] = []
        self._wall: Dict[str, float] = {}

    @contextlib.contextmanager
    def stage(self, name: str):
        t0 = time.perf_counter()
        if self.enabled:
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.nvtx.range_push(name)
            e0.record()
        try:
            yield
        finally:
            if self.enabled:
                e1.record()
                torch.cuda.nvtx.range_pop()
                self._events.append((name, e0, e1))
            self._wall[name] = self._wall.get(name, 0.0) + (time.perf_counter() - t0) * 1e3

    def result(self) -> Dict[str, float]:
        """{stage: milliseconds}; device time where available, else wall time."""
        out: Dict[str, float] = {}
        if self.enabled:
            torch.cuda.synchronize()
            for name, e0, e1 in self._events:
                out[name] = out.get(name, 0.0) + e0.elapsed_time(e1)
        for name, ms in self._wall.items():
            out.setdefault(name, ms)
            out[name + "_wall"] = ms
        return out


class Watchdog:
    """`with Watchdog(seconds, what):` raises TimeoutError in the caller if the block does not finish
    (checked when the block exits or polls `check()`), and records it for the supervisor."""

    def __init__(self, seconds: float, what: str = "device call"):
        self.seconds, self.what = seconds, what
        self.expired = threading.Event()
        self._timer = None

    def __enter__(self):
        self._timer = threading.Timer(self.seconds, self.expired.set)
        self._timer.daemon = True
        self._timer.start()
        return self

    def check(self):
        if self.expired.is_set():
            raise TimeoutError(f"{self.what} exceeded {self.seconds} s")

    def __exit__(self, exc_type, exc, tb):
        self._timer.cancel()
        if exc_type is None:
            self.check()
        return False
