"""Tracing / metrics helpers (the reference has logging only): NVTX ranges, CUDA-event step
timers and a JSON-lines metrics sink, all optional and zero-cost when disabled."""
from __future__ import annotations

import contextlib
import json
import os
import time
from typing import Dict, Optional

import torch


@contextlib.contextmanager
def nvtx_range(name: str):
    """NVTX range visible in Nsight tools (no-op without CUDA)."""
    on = torch.cuda.is_available()
    if on:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if on:
            torch.cuda.nvtx.range_pop()


class StepTimer:
    """Device time of code regions measured with CUDA events on the current stream; results are read
    lazily (``summary()`` synchronises once) so timing never stalls the training loop."""

    def __init__(self, enabled: bool = True) -> None:
        self.enabled = enabled and torch.cuda.is_available()
        self._pending = []
        self._totals: Dict[str, float] = {}
        self._counts: Dict[str, int] = {}

    @contextlib.contextmanager
    def region(self, name: str):
        if not self.enabled:
            yield
            return
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        try:
            yield
        finally:
            e.record()
            self._pending.append((name, s, e))

    def summary(self) -> Dict[str, float]:
        if self.enabled and self._pending:
            torch.cuda.synchronize()
            for name, s, e in self._pending:
                self._totals[name] = self._totals.get(name, 0.0) + s.elapsed_time(e)
                self._counts[name] = self._counts.get(name, 0) + 1
            self._pending.clear()
        return {k: self._totals[k] / self._counts[k] for k in self._totals}


class JsonlMetrics:
    """Append-only JSON-lines metrics file (one object per event: step, loss, docs/s, ...)."""

    def __init__(self, path: Optional[str]) -> None:
        self.path = path
        self._f = None
        if path:
            os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
            self._f = open(path, "a", buffering=1)

    def log(self, **fields) -> None:
        if self._f is not None:
            fields.setdefault("ts", time.time())
            self._f.write(json.dumps(fields) + "\n")

    def close(self) -> None:
        if self._f is not None:
            self._f.close()
            self._f = None
