"""A whole federation on ONE GPU: N logical clients, the fused aggregation kernel emulating every rank.

``csrc/fabric.cu`` can run all ranks of the federation as co-resident slices of a single grid (one
``GfFabricRank`` context per ``blockIdx.y``); the peer pointers then are ordinary pointers into N arenas of the
same device.  The kernels, the flag protocol and the arithmetic are exactly those of the multi-GPU run (stand-alone
aggregation path), so this is both how the protocol is tested without an NVSwitch box and a cheap way to study the
*statistical* behaviour of a federation — topic quality against the number of clients, ``agg="grad"`` against
``agg="param"``, subset sharing — at the cost of one GPU.

Semantics: src/federation/server.py:430-521 (synchronous per-minibatch aggregation weighted by corpus size) and
federated_model.py:98-131 of the reference.
"""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..parallel.fabric import EmulatedArena, FusedAggregator


class EmulatedFederation:
    def __init__(self, trainers: Sequence, datasets: Sequence, agg: str = "grad",
                 grads_to_share: Optional[Sequence[str]] = None) -> None:
        """``trainers``: one fused-backend ``AVITM`` / ``CTM`` per logical client (same architecture); ``datasets``:
        their corpora.  Every client starts from client 0's weights (the reference's intent, server.py:292-301)."""
        from ..engine.fused import FusedEngine
        from .trainer import FederatedTrainer

        if agg not in ("grad", "param"):
            raise ValueError("agg must be 'grad' or 'param'")
        self.trainers, self.agg = list(trainers), agg
        n = self.world = len(self.trainers)
        self.loaders = [t.make_loader(d, shuffle=True) for t, d in zip(self.trainers, datasets)]
        sizes = [float(len(d)) for d in datasets]
        self.weights = [s / sum(sizes) for s in sizes]
        arena = EmulatedArena(n)
        self.engines: List[FusedEngine] = []
        for r, (t, ld) in enumerate(zip(self.trainers, self.loaders)):
            if t.backend != "fused":
                raise RuntimeError("EmulatedFederation drives the fused engine")
            t.engine = FusedEngine(t.model, max_batch=ld.batch_size, max_nnz=max(ld.max_nnz, 1), lr=t.lr,
                                   betas=(t.momentum, 0.99), kl_weight=float(t.weights["beta"]), seed=t.seed + r,
                                   device=t.device, arena=arena if r == 0 else arena.view_as(r), solver=t.solver)
            if r == 0:
                arena = t.engine.flat.arena
            self.engines.append(t.engine)
        e0 = self.engines[0]
        for e in self.engines[1:]:
            e.flat.P.copy_(e0.flat.P)
            e.flat.S.copy_(e0.flat.S)
            e.sync_shadows()
        fl = e0.flat
        planner = FederatedTrainer.__new__(FederatedTrainer)
        planner.grads_to_share = list(grads_to_share) if grads_to_share else None
        planner.model, planner.agg, planner.world, planner.arena = self.trainers[0], agg, 1, arena
        ranges, s_mask, self.beta_shared = planner._share_plan(fl)
        t0 = self.trainers[0]
        self.aggregator = FusedAggregator(world=n, weights=self.weights, flat_n=fl.beta_off, beta_off=fl.beta_off,
                                          s_n=fl.s_n, K=e0.K, V=e0.V, Kp=e0.Kp,
                                          ldb=(e0.beta_bf.stride(0) if e0.is_prod else 0), lr=t0.lr,
                                          betas=e0.betas, eps=e0.adam_eps, solver=e0.solver, ranges=ranges, s_mask=s_mask,
                                          beta_shared=self.beta_shared, local_gscale=1.0)
        for r, e in enumerate(self.engines):
            self.aggregator.add_rank(arena.view_as(r) if r else arena, r, M=e.flat.M, V2=e.flat.V2, has_shadows=e.is_prod)
        self._iters = [iter(ld) for ld in self.loaders]
        self.global_step = 0

    def _next(self, r: int):
        try:
            return next(self._iters[r])
        except StopIteration:
            self._iters[r] = iter(self.loaders[r])
            return next(self._iters[r])

    @property
    def steps_per_epoch(self) -> int:
        return max(len(ld) for ld in self.loaders)

    def step(self) -> List[torch.Tensor]:
        """One synchronous global step of every client + the aggregation.  Returns the clients' summed losses."""
        self.global_step += 1
        losses = []
        for r, e in enumerate(self.engines):
            mb = self._next(r)
            if self.agg == "grad":
                # the emulation shares one argument block, so unshared tensors use a unit local gradient scale:
                # fold the client weight into shared tensors only when everything is shared
                losses.append(e.forward_backward(mb.bow, mb.x_ctx, training=True,
                                                 gscale=self.weights[r] if self.aggregator.args.n_ranges == 0 else 1.0).clone())
                if not self.beta_shared:
                    e.step_count = self.global_step - 1
                    e.optimizer_step_beta(gscale=1.0)
            else:
                losses.append(e.forward_backward(mb.bow, mb.x_ctx, training=True).clone())
                e.optimizer_step()
        self.aggregator.launch(mode=self.agg, step=self.global_step)
        for e in self.engines:
            e.step_count = self.global_step
            e._refresh_ctm_shadows()
            e._refresh_ctx_shadows()
        return losses

    def fit(self, epochs: int, log_every: int = 0, logger=None) -> None:
        for t in self.trainers:
            t.model.train()
        total = epochs * self.steps_per_epoch
        for s in range(total):
            losses = self.step()
            if log_every and logger is not None and (s + 1) % log_every == 0:
                logger.info("global step %d/%d  mean loss/doc %.3f", s + 1, total,
                            sum(float(l.item()) for l in losses) / sum(ld.batch_size for ld in self.loaders))
        torch.cuda.synchronize()
        for t in self.trainers:
            t.best_components = t.model.beta
            t.engine.flush_counters()
