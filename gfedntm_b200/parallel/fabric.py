"""NVLink / NVSwitch fabric layer: a symmetric arena shared by the federated clients of one box and
the launcher of the fused aggregation kernel (``csrc/fabric.cu``).

One process per GPU.  ``torch.distributed`` (NCCL backend) is used only for the plumbing:
rendezvous of the symmetric allocation through ``torch.distributed._symmetric_memory`` (CUDA VMM
handles exchanged between the processes, every peer's arena mapped into every process) and host
barriers.  The per-step data path is the kernel, which loads/stores peer memory directly.

Three arena flavours:

* ``SymmetricArena``  — real multi-GPU (world > 1).
* ``LocalArena``      — world == 1 (plain device memory).
* ``EmulatedArena``   — N logical ranks on ONE device for tests: every "peer" buffer is a separate
  allocation on the same GPU and the kernel runs all ranks as co-resident blocks of one grid.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Dict, List, Optional, Sequence

import torch

from ..ops import _lib
from ..ops._lib import GF_FED_FLAG_BYTES, GF_FED_MAX_RANGES, GF_MAX_RANKS, GfFedArgs, check


def _align(n: int, a: int = 256) -> int:
    return (n + a - 1) // a * a


class _ArenaBase:
    """Bump allocator over one byte buffer; remembers the offset of every named block."""

    def __init__(self) -> None:
        self.offsets: Dict[str, int] = {}
        self.sizes: Dict[str, int] = {}
        self._cursor = 0

    def reserve(self, name: str, nbytes: int) -> None:
        self.offsets[name] = self._cursor
        self.sizes[name] = nbytes
        self._cursor = _align(self._cursor + nbytes)

    @property
    def nbytes(self) -> int:
        return max(self._cursor, 256)

    def _view(self, buf: torch.Tensor, name: str, dtype: torch.dtype) -> torch.Tensor:
        off, n = self.offsets[name], self.sizes[name]
        return buf[off:off + n].view(dtype)


class LocalArena(_ArenaBase):
    world, rank = 1, 0

    def commit(self, device) -> None:
        self.buf = torch.zeros(self.nbytes, dtype=torch.uint8, device=device)
        self.base_ptrs = [self.buf.data_ptr()]

    def tensor(self, name: str, dtype: torch.dtype) -> torch.Tensor:
        return self._view(self.buf, name, dtype)

    def peer_ptr(self, rank: int, name: str) -> int:
        return self.base_ptrs[rank] + self.offsets[name]

    def barrier(self) -> None:
        torch.cuda.synchronize()


class SymmetricArena(_ArenaBase):
    """Symmetric allocation across the ranks of a ``torch.distributed`` process group."""

    def __init__(self, group=None) -> None:
        super().__init__()
        import torch.distributed as dist

        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)

    def commit(self, device) -> None:
        import torch.distributed as dist
        import torch.distributed._symmetric_memory as symm_mem

        self.buf = symm_mem.empty(self.nbytes, dtype=torch.uint8, device=device)
        self.buf.zero_()
        self.handle = symm_mem.rendezvous(self.buf, self.group)
        self.base_ptrs = [int(p) for p in self.handle.buffer_ptrs]
        self.multicast_ptr = int(getattr(self.handle, "multicast_ptr", 0) or 0)
        torch.cuda.synchronize()
        dist.barrier(self.group)

    def tensor(self, name: str, dtype: torch.dtype) -> torch.Tensor:
        return self._view(self.buf, name, dtype)

    def peer_ptr(self, rank: int, name: str) -> int:
        return self.base_ptrs[rank] + self.offsets[name]

    def barrier(self) -> None:
        import torch.distributed as dist

        torch.cuda.synchronize()
        dist.barrier(self.group)


class EmulatedArena(_ArenaBase):
    """``world`` logical ranks on one device.  ``view_as(rank)`` returns that rank's arena."""

    def __init__(self, world: int) -> None:
        super().__init__()
        self.world = world
        self.rank = 0
        self._shared: Optional["EmulatedArena"] = None

    def commit(self, device) -> None:
        self.bufs = [torch.zeros(self.nbytes, dtype=torch.uint8, device=device) for _ in range(self.world)]
        self.base_ptrs = [b.data_ptr() for b in self.bufs]

    def view_as(self, rank: int) -> "EmulatedArena":
        v = EmulatedArena(self.world)
        v.offsets, v.sizes, v._cursor = self.offsets, self.sizes, self._cursor
        v.bufs, v.base_ptrs, v.rank = self.bufs, self.base_ptrs, rank
        return v

    def tensor(self, name: str, dtype: torch.dtype) -> torch.Tensor:
        return self._view(self.bufs[self.rank], name, dtype)

    def peer_ptr(self, rank: int, name: str) -> int:
        return self.base_ptrs[rank] + self.offsets[name]

    def barrier(self) -> None:
        torch.cuda.synchronize()


FLAG_BYTES = GF_FED_FLAG_BYTES  # uint32 [GF_FED_FLAG_SLOTS][GF_MAX_RANKS] per rank (include/gfedntm_kernels.h)


def fill_rank_ctx(ctx, arena, rank: int, *, M: torch.Tensor, V2: torch.Tensor, done_counter: torch.Tensor,
                  has_shadows: bool, agg_counter: Optional[torch.Tensor] = None, ld_gb: int = 0) -> None:
    """Peer pointer tables of one rank (names: P, G, S, beta_bf, betaT_bf, Gb, flags)."""
    ctx.rank = rank
    has_gb = has_shadows and "Gb" in arena.offsets
    for c in range(arena.world):
        ctx.Gb[c] = arena.peer_ptr(c, "Gb") if has_gb else None
    ctx.ld_gb = ld_gb if has_gb else 0
    ctx.agg_counter = agg_counter.data_ptr() if agg_counter is not None else None
    has_ar = "AR" in arena.offsets
    for c in range(arena.world):
        ctx.AR[c] = arena.peer_ptr(c, "AR") if has_ar else None
    ctx.ar_cap = (arena.sizes["AR"] // (4 * arena.world)) if has_ar else 0
    for c in range(arena.world):
        ctx.G[c] = arena.peer_ptr(c, "G")
        ctx.P[c] = arena.peer_ptr(c, "P")
        ctx.S[c] = arena.peer_ptr(c, "S")
        ctx.flags[c] = arena.peer_ptr(c, "flags")
        ctx.beta_bf[c] = arena.peer_ptr(c, "beta_bf") if has_shadows else None
        ctx.betaT_bf[c] = arena.peer_ptr(c, "betaT_bf") if has_shadows else None
    ctx.M, ctx.V2 = M.data_ptr(), V2.data_ptr()
    ctx.done_counter = done_counter.data_ptr()
    # NVLS: one multimem.st per element instead of one P2P store per rank
    mc = int(getattr(arena, "multicast_ptr", 0) or 0)
    # GFEDNTM_NVLS=1 / =0 forces it; by default it is used from 8 ranks on, where it was measured (same box,
    # back to back: 3.41 ms/step with 32 aggregation CTAs against 3.50 with P2P or multicast stores and 40)
    env = os.environ.get("GFEDNTM_NVLS")
    use_mc = mc != 0 and (env == "1" or (env not in ("0", "1") and arena.world >= 8))
    ctx.P_mc = (mc + arena.offsets["P"]) if use_mc else None
    ctx.S_mc = (mc + arena.offsets["S"]) if use_mc else None
    ctx.beta_bf_mc = (mc + arena.offsets["beta_bf"]) if (use_mc and has_shadows) else None
    ctx.betaT_bf_mc = (mc + arena.offsets["betaT_bf"]) if (use_mc and has_shadows) else None


class FusedAggregator:
    """Launches ``gf_fed_aggregate`` for one rank (or for all emulated ranks at once)."""

    def __init__(self, *, world: int, weights: Sequence[float], flat_n: int, beta_off: int, s_n: int, K: int,
                 V: int, Kp: int, ldb: int, lr: float, betas, eps: float = 1e-8, wd: float = 0.0,
                 ranges: Optional[Sequence[tuple]] = None, s_mask: int = 63, beta_shared: bool = True,
                 local_gscale: float = 1.0, solver: int = 0) -> None:
        """``ranges``: optional ``(offset, length, shared)`` pieces covering the flat region ``[0, flat_n)``
        (grads_to_share of the reference, federated_model.py:98-131): unshared pieces stay client-local.
        ``s_mask``: bit per BatchNorm statistics block; ``beta_shared``: whether beta takes part."""
        assert 1 <= world <= GF_MAX_RANKS
        self.args = GfFedArgs()
        a = self.args
        a.n_ranges = 0
        a.s_mask, a.beta_shared, a.local_gscale = int(s_mask), int(beta_shared), float(local_gscale)
        a.solver = int(solver)  # GfSolver of the global update (b1 = momentum, b2 = rho / alpha, see opt_math.cuh)
        if ranges is not None:
            merged: List[list] = []
            for off, ln, sh in ranges:
                if ln <= 0:
                    continue
                assert off % 4 == 0 and ln % 4 == 0, "ranges must be 16-byte aligned"
                kind = int(sh)  # 0 client-local, 1 aggregated, 2 replicated (see GfFedRange)
                if merged and merged[-1][2] == kind and merged[-1][0] + merged[-1][1] == off:
                    merged[-1][1] += ln
                else:
                    merged.append([off, ln, kind])
            all_shared = all(m[2] == 1 for m in merged) and (s_mask & 63) == 63 and beta_shared
            if not all_shared:
                assert len(merged) <= GF_FED_MAX_RANGES, "too many alternating shared / unshared segments"
                a.n_ranges = len(merged)
                for i, (off, ln, sh) in enumerate(merged):
                    a.ranges[i].off, a.ranges[i].len, a.ranges[i].shared = off, ln, int(sh)
        a.world, a.flat_n, a.beta_off, a.s_n = world, flat_n, beta_off, s_n
        a.K, a.V, a.Kp, a.ldb = K, V, Kp, ldb
        for c in range(world):
            a.w[c] = float(weights[c])
        a.lr, a.b1, a.b2, a.eps, a.wd = lr, betas[0], betas[1], eps, wd
        a.n_rank_ctx = 0
        self.epoch = 0
        self._keep: List[torch.Tensor] = []

    def add_rank(self, arena, rank: int, *, M, V2, has_shadows: bool, ld_gb: int = 0) -> None:
        i = self.args.n_rank_ctx
        done = torch.zeros(2, dtype=torch.int32, device=M.device)  # [0] launch-completion, [1] word-tile counter
        self._keep.append(done)
        fill_rank_ctx(self.args.ranks[i], arena, rank, M=M, V2=V2, done_counter=done[0:1], has_shadows=has_shadows,
                      agg_counter=done[1:2], ld_gb=ld_gb)
        self.args.n_rank_ctx = i + 1
        self._arena = arena

    def set_lr(self, lr: float, betas) -> None:
        self.args.lr, self.args.b1, self.args.b2 = lr, betas[0], betas[1]

    # ---- overlapped beta aggregation (grad mode): see csrc/dec_sweep.cu, GF_ROW_GRAD2 with `fed`
    def overlap_handle(self, step_ctr: torch.Tensor, ctas: int) -> "BetaOverlap":
        """Uploads the aggregation arguments once (they are constant across steps: the epoch and the Adam
        step are read from ``step_ctr`` on the device) and returns the handle the decoder uses."""
        assert self.args.n_rank_ctx == 1, "the overlapped path runs one rank per process"
        a = self.args
        assert a.ranks[0].Gb[0], "the arena has no bf16 gradient buffer (Gb)"
        a.mode, a.epoch, a.step, a.parts = 0, 0, 0, 2
        a.step_ctr = step_ctr.data_ptr()
        raw = bytes(memoryview(a).cast("B"))
        dev = torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(step_ctr.device)
        self._keep.append(dev)
        # tensor maps of every rank's bf16 beta gradient (host memory: they travel as a kernel parameter)
        tm = (C.c_ubyte * (128 * (GF_MAX_RANKS + 3) + 64))()  # GF_FED_TMAPS
        tm_ptr = (C.addressof(tm) + 63) // 64 * 64
        check(_lib.load().gf_fed_make_tmaps(C.byref(a), tm_ptr), "gf_fed_make_tmaps")
        self._keep.append(tm)
        gb = self._arena.tensor("Gb", torch.bfloat16).view(a.K, a.ranks[0].ld_gb)
        return BetaOverlap(self, dev, step_ctr, ctas, tm_ptr, gb)

    def allreduce_small(self, x: torch.Tensor, step_ctr: torch.Tensor) -> None:
        """In-place sum of a small fp32 buffer over the ranks (one-shot all-reduce through peer memory); the
        flag epoch is the device step counter, so the launch replays from a CUDA graph."""
        a = self.args
        a.step_ctr = step_ctr.data_ptr()
        n = x.numel()
        assert n % 4 == 0 and x.is_contiguous()
        check(_lib.load().gf_fed_allreduce_small(C.byref(a), x.data_ptr(), n, torch.cuda.current_stream().cuda_stream),
              "gf_fed_allreduce_small")

    def gather_beta(self) -> None:
        """Completes this rank's fp32 copy of beta from the owners' master slices (overlapped path)."""
        check(_lib.load().gf_fed_gather_beta(C.byref(self.args), torch.cuda.current_stream().cuda_stream),
              "gf_fed_gather_beta")

    def launch(self, *, mode: str, step: int, step_ctr=None, parts: int = 3) -> None:
        """``mode``: 'grad' (weighted gradients + global sharded Adam) or 'param' (weighted mean of the
        parameters).  Every rank must call this once per global step.  With ``step_ctr`` (device counters
        of the engine) the epoch flag and the Adam step are read on the device, so the launch can be
        replayed from a CUDA graph."""
        self.epoch += 1
        a = self.args
        a.mode = 0 if mode == "grad" else 1
        a.epoch = self.epoch
        a.step = step
        a.step_ctr = step_ctr.data_ptr() if step_ctr is not None else None
        a.parts = parts
        lib = _lib.load()
        check(lib.gf_fed_aggregate(C.byref(a), torch.cuda.current_stream().cuda_stream), "gf_fed_aggregate")


class BetaOverlap:
    """Handle passed down to the decoder: raises this rank's "beta gradient ready" flag on the peers and
    carries the device copy of the aggregation arguments for the sweep kernel's aggregation CTAs."""

    PHASE_BETA_READY = 2

    def __init__(self, agg: FusedAggregator, dev_args: torch.Tensor, step_ctr: torch.Tensor, ctas: int,
                 tmaps_ptr: int, gb: torch.Tensor) -> None:
        self.agg, self._dev, self.step_ctr, self.ctas = agg, dev_args, step_ctr, int(ctas)
        self.dev_args = dev_args.data_ptr()
        self.tmaps_ptr = tmaps_ptr   # host pointer: GF_MAX_RANKS CUtensorMap (gf_fed_make_tmaps)
        self.gb = gb                 # [K, ld_gb] bf16: this rank's beta gradient in the exchange format

    def signal_beta_ready(self) -> None:
        a = self.agg.args
        a.step_ctr = self.step_ctr.data_ptr()
        check(_lib.load().gf_fed_signal(C.byref(a), self.PHASE_BETA_READY, torch.cuda.current_stream().cuda_stream),
              "gf_fed_signal")
