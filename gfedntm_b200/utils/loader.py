"""Minibatch loader that feeds CSR minibatches (plus contextual embeddings / labels) to the device.

The corpus stays on the host as one CSR matrix.  Each minibatch is assembled by the native helper
``gf_host_gather_csr`` straight into pinned staging buffers and copied with one async H2D copy
per array.  On CUDA a background thread assembles and uploads the following batches (three staging slots
rotate) while the current step runs, so neither the host gather nor the H2D copy sits on the step's
critical path even when the caller synchronises on every step's loss.

Ref: ``DataLoader(train_data, batch_size, shuffle=True)`` in federated_model.py:82-86 and
avitm.py:371-375 (dense rows, default collate, ``.cuda()`` per batch).
"""
from __future__ import annotations

import queue
import threading
from dataclasses import dataclass
from typing import Iterator, Optional

import numpy as np
import torch

from ..ops.decoder import CsrBatch


@dataclass
class Minibatch:
    bow: CsrBatch
    x_ctx: Optional[torch.Tensor]
    labels: Optional[torch.Tensor]
    index: np.ndarray          # corpus row ids of the documents in this batch
    h2d_bytes: int = 0

    @property
    def size(self) -> int:
        return self.bow.B


def _native():
    try:
        from ..ops import _lib

        return _lib.load()
    except Exception:  # library not built (pure-CPU install): numpy fallback below
        return None


class _Slot:
    def __init__(self, max_B: int, max_nnz: int, C_ctx: int, L: int, device: torch.device) -> None:
        pin = device.type == "cuda"
        self.indptr_h = torch.empty(max_B + 1, dtype=torch.int32, pin_memory=pin)
        self.indices_h = torch.empty(max(max_nnz, 1), dtype=torch.int32, pin_memory=pin)
        self.counts_h = torch.empty(max(max_nnz, 1), dtype=torch.float32, pin_memory=pin)
        self.ctx_h = torch.empty(max_B, C_ctx, dtype=torch.float32, pin_memory=pin) if C_ctx else None
        self.lab_h = torch.empty(max_B, L, dtype=torch.float32, pin_memory=pin) if L else None
        if pin:
            self.indptr_d = torch.empty(max_B + 1, dtype=torch.int32, device=device)
            self.indices_d = torch.empty(max(max_nnz, 1), dtype=torch.int32, device=device)
            self.counts_d = torch.empty(max(max_nnz, 1), dtype=torch.float32, device=device)
            self.ctx_d = torch.empty(max_B, C_ctx, dtype=torch.float32, device=device) if C_ctx else None
            self.lab_d = torch.empty(max_B, L, dtype=torch.float32, device=device) if L else None
        self.ready = torch.cuda.Event() if pin else None
        self.consumed = None                      # CUDA event: every kernel reading this slot has been queued
        self.free = threading.Event()             # host side: `consumed` is valid (or the slot is unused)
        self.free.set()
        self.in_flight = False


class CsrMinibatchLoader:
    """Iterates over a ``BOWDataset`` / ``CTMDataset`` in (optionally shuffled) CSR minibatches."""

    def __init__(self, dataset, batch_size: int, shuffle: bool = True, drop_last: bool = False,
                 device=None, seed: int = 0, n_threads: int = 4, prefetch: Optional[bool] = None) -> None:
        self.dataset = dataset
        self.batch_size = int(batch_size)
        self.shuffle, self.drop_last = shuffle, drop_last
        self.device = torch.device(device) if device is not None else torch.device("cpu")
        self.rng = np.random.default_rng(seed)
        self.n_threads = n_threads
        X = dataset.X
        self.indptr = np.ascontiguousarray(X.indptr, dtype=np.int32)
        self.indices = np.ascontiguousarray(X.indices, dtype=np.int32)
        self.data = np.ascontiguousarray(X.data, dtype=np.float32)
        self.N, self.V = X.shape
        lens = np.diff(self.indptr)
        top = np.sort(lens)[::-1][: self.batch_size]
        self.max_nnz = int(top.sum()) if len(top) else 0
        self.ctx = getattr(dataset, "X_contextual", None)
        self.labels = getattr(dataset, "labels", None)
        C_ctx = self.ctx.shape[1] if self.ctx is not None else 0
        L = self.labels.shape[1] if self.labels is not None else 0
        if prefetch is None:
            # the hand-off to the staging thread costs ~30 us per batch: only worth it when assembling a
            # batch takes longer than that (measured: B=64 217k -> 199k docs/s, B=4096 1.02M -> 1.09M)
            prefetch = self.max_nnz >= 65536
        self.prefetch = bool(prefetch) and self.device.type == "cuda"
        self.slots = [_Slot(self.batch_size, self.max_nnz, C_ctx, L, self.device)
                      for _ in range(3 if self.prefetch else 2)]
        self.lib = _native()
        self.copy_stream = torch.cuda.Stream(self.device) if self.device.type == "cuda" else None

    def __len__(self) -> int:
        if self.drop_last:
            return self.N // self.batch_size
        return (self.N + self.batch_size - 1) // self.batch_size

    # -- host assembly + H2D of one batch into a slot
    def _stage(self, slot: _Slot, rows: np.ndarray) -> Minibatch:
        n = len(rows)
        if slot.in_flight:
            slot.ready.synchronize()  # the previous H2D out of these pinned buffers must be done
        rows64 = np.ascontiguousarray(rows, dtype=np.int64)
        ip, ix, ct = slot.indptr_h.numpy(), slot.indices_h.numpy(), slot.counts_h.numpy()
        if self.lib is not None:
            rc = self.lib.gf_host_gather_csr(self.indptr.ctypes.data, self.indices.ctypes.data,
                                             self.data.ctypes.data, rows64.ctypes.data, n, ip.ctypes.data,
                                             ix.ctypes.data, ct.ctypes.data, len(ix), self.n_threads)
            if rc != 0:
                raise RuntimeError("gf_host_gather_csr failed (staging buffer too small?)")
        else:
            lens = self.indptr[rows64 + 1] - self.indptr[rows64]
            ip[0] = 0
            np.cumsum(lens, out=ip[1:n + 1])
            for i, r in enumerate(rows64):
                s, e = self.indptr[r], self.indptr[r + 1]
                ix[ip[i]:ip[i + 1]] = self.indices[s:e]
                ct[ip[i]:ip[i + 1]] = self.data[s:e]
        nnz = int(ip[n])
        if self.ctx is not None:
            if self.lib is not None:
                self.lib.gf_host_gather_rows(self.ctx.ctypes.data, self.ctx.shape[1], rows64.ctypes.data, n,
                                             slot.ctx_h.numpy().ctypes.data)
            else:
                slot.ctx_h.numpy()[:n] = self.ctx[rows64]
        if self.labels is not None:
            slot.lab_h.numpy()[:n] = self.dataset.label_rows(rows64)
        h2d = 4 * (n + 1) + 8 * nnz
        if self.device.type == "cuda":
            with torch.cuda.stream(self.copy_stream):
                slot.indptr_d[: n + 1].copy_(slot.indptr_h[: n + 1], non_blocking=True)
                slot.indices_d[:nnz].copy_(slot.indices_h[:nnz], non_blocking=True)
                slot.counts_d[:nnz].copy_(slot.counts_h[:nnz], non_blocking=True)
                if slot.ctx_h is not None:
                    slot.ctx_d[:n].copy_(slot.ctx_h[:n], non_blocking=True)
                    h2d += 4 * n * slot.ctx_h.shape[1]
                if slot.lab_h is not None:
                    slot.lab_d[:n].copy_(slot.lab_h[:n], non_blocking=True)
                    h2d += 4 * n * slot.lab_h.shape[1]
                slot.ready.record(self.copy_stream)
                slot.in_flight = True
            bow = CsrBatch(slot.indptr_d[: n + 1], slot.indices_d[:nnz], slot.counts_d[:nnz], n, self.V)
            ctx = slot.ctx_d[:n] if slot.ctx_h is not None else None
            lab = slot.lab_d[:n] if slot.lab_h is not None else None
        else:
            bow = CsrBatch(slot.indptr_h[: n + 1].clone(), slot.indices_h[:nnz].clone(),
                           slot.counts_h[:nnz].clone(), n, self.V)
            ctx = slot.ctx_h[:n].clone() if slot.ctx_h is not None else None
            lab = slot.lab_h[:n].clone() if slot.lab_h is not None else None
        return Minibatch(bow, ctx, lab, rows64, h2d)

    def _epoch_rows(self) -> Iterator[np.ndarray]:
        order = self.rng.permutation(self.N) if self.shuffle else np.arange(self.N)
        starts = list(range(0, self.N, self.batch_size))
        if self.drop_last and starts and self.N - starts[-1] < self.batch_size:
            starts.pop()
        for s in starts:
            yield order[s:s + self.batch_size]

    def _rows(self, endless: bool) -> Iterator[np.ndarray]:
        while True:
            yield from self._epoch_rows()
            if not endless:
                return

    def __iter__(self) -> Iterator[Minibatch]:
        return self._iterate(self._rows(False))

    def cycle(self) -> Iterator[Minibatch]:
        """Endless iteration, a new permutation per epoch: the staging pipeline (worker thread, H2D copies two batches
        ahead) keeps running across epoch boundaries instead of draining and restarting at every epoch."""
        return self._iterate(self._rows(True))

    def _iterate(self, rows_iter: Iterator[np.ndarray]) -> Iterator[Minibatch]:
        if self.prefetch:
            yield from self._iter_prefetch(rows_iter)
            return
        pending = None
        for i, rows in enumerate(rows_iter):
            slot = self.slots[i & 1]
            if self.device.type == "cuda":
                # the slot's device buffers were last read by the step two batches ago
                self.copy_stream.wait_stream(torch.cuda.current_stream(self.device))
            mb = self._stage(slot, rows)
            if pending is not None:
                yield pending
            if self.device.type == "cuda":
                torch.cuda.current_stream(self.device).wait_event(slot.ready)
            pending = mb
        if pending is not None:
            yield pending

    # -- background staging: the worker is at most two batches ahead (one queued, one being assembled)
    def _iter_prefetch(self, rows_iter: Iterator[np.ndarray]) -> Iterator[Minibatch]:
        q: "queue.Queue" = queue.Queue(maxsize=1)
        stop = threading.Event()
        dev = self.device

        def worker() -> None:
            try:
                torch.cuda.set_device(dev)
                for i, rows in enumerate(rows_iter):
                    slot = self.slots[i % 3]
                    while not slot.free.wait(0.05):
                        if stop.is_set():
                            return
                    if slot.consumed is not None:
                        self.copy_stream.wait_event(slot.consumed)  # the step that read this slot is queued
                    mb = self._stage(slot, rows)
                    while True:
                        if stop.is_set():
                            return
                        try:
                            q.put((mb, slot), timeout=0.05)
                            break
                        except queue.Full:
                            continue
                q.put(None)
            except BaseException as ex:  # surface staging errors in the consumer
                q.put(ex)

        t = threading.Thread(target=worker, name="gfedntm-loader", daemon=True)
        t.start()
        prev = None
        try:
            while True:
                item = q.get()
                if item is None:
                    break
                if isinstance(item, BaseException):
                    raise item
                mb, slot = item
                cur = torch.cuda.current_stream(dev)
                if prev is not None:  # everything that reads the previous batch has been queued by now
                    prev.consumed = torch.cuda.Event()
                    prev.consumed.record(cur)
                    prev.free.set()
                slot.free.clear()
                cur.wait_event(slot.ready)
                prev = slot
                yield mb
        finally:
            stop.set()
            if prev is not None:
                prev.consumed = torch.cuda.Event()
                prev.consumed.record(torch.cuda.current_stream(dev))
                prev.free.set()
            for sl in self.slots:
                sl.free.set()
            t.join(timeout=5)
