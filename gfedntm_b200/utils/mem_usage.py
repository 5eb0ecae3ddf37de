"""Process / GPU resource sampler (the role of the submodule's ``Mem`` helper,
src/aux_modules/topicmodeler/src/utils/mem_usage.py:10-153): polls the RSS / VMS / CPU share of a set of
processes - and, when NVML is importable, the memory and utilisation of every visible GPU - at a fixed
period and writes one CSV row per process and sample.  Implemented as a background thread with a
context-manager interface instead of a signal-driven foreground loop, so a trainer can wrap itself:

    with ResourceSampler("usage.csv", period_s=1.0, gpu=True):
        model.fit(dataset)
"""
from __future__ import annotations

import csv
import os
import threading
import time
from typing import Iterable, List, Optional

import psutil


class ResourceSampler:
    HEADER = ["t", "pid", "name", "rss", "vms", "cpu_percent"]

    def __init__(self, path: str, period_s: float = 1.0, pids: Optional[Iterable[int]] = None,
                 name_contains: Optional[Iterable[str]] = None, gpu: bool = False) -> None:
        self.path, self.period = path, float(period_s)
        self.pids = list(pids) if pids is not None else [os.getpid()]
        self.name_contains = [s.lower() for s in (name_contains or [])]
        self._stop = threading.Event()
        self._thread: Optional[threading.Thread] = None
        self._nvml = None
        if gpu:
            try:
                import pynvml

                pynvml.nvmlInit()
                self._nvml = pynvml
            except Exception:  # no driver / no NVML on this host: CPU columns only
                self._nvml = None
        self.rows_written = 0

    # -- selection of the processes to watch
    def _processes(self) -> List[psutil.Process]:
        out = []
        for pid in self.pids:
            try:
                out.append(psutil.Process(pid))
            except psutil.Error:
                pass
        if self.name_contains:
            for p in psutil.process_iter(["name"]):
                try:
                    nm = (p.info.get("name") or "").lower()
                    if any(s in nm for s in self.name_contains) and p.pid not in self.pids:
                        out.append(p)
                except psutil.Error:
                    continue
        return out

    def _gpu_columns(self) -> List[str]:
        if self._nvml is None:
            return []
        cols = []
        for i in range(self._nvml.nvmlDeviceGetCount()):
            cols += [f"gpu{i}_mem_used", f"gpu{i}_util"]
        return cols

    def _gpu_values(self) -> List[float]:
        vals: List[float] = []
        if self._nvml is None:
            return vals
        nv = self._nvml
        for i in range(nv.nvmlDeviceGetCount()):
            h = nv.nvmlDeviceGetHandleByIndex(i)
            vals += [float(nv.nvmlDeviceGetMemoryInfo(h).used), float(nv.nvmlDeviceGetUtilizationRates(h).gpu)]
        return vals

    def sample_once(self, writer, t0: float) -> None:
        gpu_vals = self._gpu_values()
        for p in self._processes():
            try:
                with p.oneshot():
                    mem = p.memory_info()
                    row = [round(time.time() - t0, 3), p.pid, p.name(), mem.rss, mem.vms, p.cpu_percent(None)]
            except psutil.Error:
                continue
            writer.writerow(row + gpu_vals)
            self.rows_written += 1

    def _run(self) -> None:
        t0 = time.time()
        with open(self.path, "w", newline="") as f:
            w = csv.writer(f)
            w.writerow(self.HEADER + self._gpu_columns())
            while not self._stop.is_set():
                self.sample_once(w, t0)
                f.flush()
                self._stop.wait(self.period)
            self.sample_once(w, t0)

    def start(self) -> "ResourceSampler":
        self._thread = threading.Thread(target=self._run, name="gfedntm-resource-sampler", daemon=True)
        self._thread.start()
        return self

    def stop(self) -> None:
        self._stop.set()
        if self._thread is not None:
            self._thread.join(timeout=5 * self.period + 5)

    def __enter__(self) -> "ResourceSampler":
        return self.start()

    def __exit__(self, *exc) -> None:
        self.stop()
