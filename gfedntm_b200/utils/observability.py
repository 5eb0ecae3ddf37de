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


class PrometheusExporter:
    """Optional metrics endpoint of the federation server (``prometheus_client`` is pinned but unused by the
    reference).  ``PrometheusExporter(port)`` starts the HTTP endpoint; with ``port=None`` (or
    without the package) every method is a no-op, so callers never need to branch."""

    def __init__(self, port: Optional[int] = None, registry=None) -> None:
        self.enabled = False
        if port is None:
            return
        try:
            import prometheus_client as pc
        except Exception:  # package absent: stay disabled
            return
        self._pc = pc
        self.registry = registry if registry is not None else pc.CollectorRegistry()
        self.iterations = pc.Counter("gfedntm_global_iterations_total", "completed global iterations",
                                     registry=self.registry)
        self.clients = pc.Gauge("gfedntm_active_clients", "clients that answered the last pull",
                                registry=self.registry)
        self.iter_seconds = pc.Histogram("gfedntm_iteration_seconds", "wall time of one global iteration",
                                         buckets=(.005, .01, .025, .05, .1, .25, .5, 1, 2.5, 5, 10),
                                         registry=self.registry)
        self.payload_bytes = pc.Counter("gfedntm_payload_bytes_total", "tensor bytes pulled from the clients",
                                        registry=self.registry)
        self.failures = pc.Counter("gfedntm_rpc_failures_total", "client RPCs that failed after all retries",
                                   registry=self.registry)
        if port > 0:
            pc.start_http_server(port, registry=self.registry)
        self.enabled = True

    def observe_iteration(self, seconds: float, n_clients: int, payload_bytes: int = 0, failures: int = 0) -> None:
        if not self.enabled:
            return
        self.iterations.inc()
        self.clients.set(n_clients)
        self.iter_seconds.observe(seconds)
        if payload_bytes:
            self.payload_bytes.inc(payload_bytes)
        if failures:
            self.failures.inc(failures)

    def scrape(self) -> str:
        """Current exposition text (tests / debugging)."""
        return self._pc.generate_latest(self.registry).decode() if self.enabled else ""
