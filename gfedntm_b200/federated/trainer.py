"""In-box federated training: one client per GPU, aggregation fused over NVLink peer memory.

This is the B200 data plane of the federation.  The reference's server pulls every client's
post-step tensors over gRPC once per minibatch, averages them on the host weighted by corpus size
and pushes the average back (src/federation/server.py:430-521).  Here every client is a rank of a
``torch.distributed`` job on one NVSwitch domain and the same synchronous per-minibatch protocol
runs as one kernel per rank (``csrc/fabric.cu``):

* ``agg="grad"``  — product path: weighted gradient sum (weights n_c / sum n folded into the
  backward pass), ONE global Adam whose state is sharded across the ranks, updated parameters and
  bf16 operand shadows pushed to every rank.
* ``agg="param"`` — reference-parity semantics: local Adam on every client, then the weighted mean
  of the parameters (exactly server.py:477-487), Adam moments stay local.

BatchNorm running statistics are averaged with the same weights in both modes.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence

import torch

from ..ops.decoder import CsrBatch
from ..parallel.fabric import FusedAggregator, LocalArena, SymmetricArena


class FederatedTrainer:
    def __init__(self, model, n_samples: int, world_size: int = 1, rank: int = 0, agg: str = "grad",
                 arena=None, weights: Optional[Sequence[float]] = None, group=None,
                 grads_to_share: Optional[Sequence[str]] = None) -> None:
        """``grads_to_share``: state-dict keys that take part in the aggregation (the reference's
        ``[federation] grads_to_share``, federated_model.py:98-131; its server default is ``prior_mean,
        prior_variance, beta``, server.py:71).  Everything else stays client-local: its own Adam step on the
        client's unweighted gradient in ``agg="grad"``, untouched by the averaging in ``agg="param"``.
        ``None`` / empty = share every parameter and BatchNorm statistic."""
        if agg not in ("grad", "param"):
            raise ValueError("agg must be 'grad' or 'param'")
        self.grads_to_share = list(grads_to_share) if grads_to_share else None
        self.model = model
        self.n_samples = int(n_samples)
        self.world, self.rank, self.agg = int(world_size), int(rank), agg
        self.arena = arena
        self.group = group
        self.weights = list(weights) if weights is not None else None
        self.aggregator: Optional[FusedAggregator] = None
        self._agg_events: List = []
        self.global_step = 0
        # observability (utils/observability.py): NVTX range per global step, CUDA-event step timer and a JSON-lines
        # sink, switched on by GFEDNTM_METRICS_JSONL=<path> (or by assigning .metrics / .timer)
        from ..utils.observability import JsonlMetrics, StepTimer

        path = os.environ.get("GFEDNTM_METRICS_JSONL")
        self.metrics = JsonlMetrics(f"{path}.rank{rank}" if path and world_size > 1 else path)
        self.timer = StepTimer(enabled=bool(path))
        self.metrics_every = int(os.environ.get("GFEDNTM_METRICS_EVERY", "100"))

    # ------------------------------------------------------------------------------------------
    def prepare(self, loader) -> None:
        """Allocates the engine inside the shared arena, agrees on the client weights and makes every
        client start from rank 0's initial model (the reference's intent, server.py:292-301)."""
        m = self.model
        if m.backend != "fused":
            raise RuntimeError("FederatedTrainer drives the fused engine; use the gRPC federation on CPU")
        from ..engine.fused import FusedEngine

        if self.arena is None:
            self.arena = SymmetricArena(self.group) if self.world > 1 else LocalArena()
        m.engine = FusedEngine(m.model, max_batch=loader.batch_size, max_nnz=max(loader.max_nnz, 1), lr=m.lr,
                               betas=(m.momentum, 0.99), kl_weight=float(m.weights["beta"]), seed=m.seed + self.rank,
                               device=m.device, arena=self.arena, solver=getattr(m, "solver", "adam"))
        eng = m.engine
        if self.world > 1 and isinstance(self.arena, SymmetricArena):
            import torch.distributed as dist

            if self.weights is None:
                n = torch.tensor([float(self.n_samples)], device=eng.device)
                alln = [torch.zeros_like(n) for _ in range(self.world)]
                dist.all_gather(alln, n, group=self.arena.group)
                tot = sum(float(x.item()) for x in alln)
                self.weights = [float(x.item()) / tot for x in alln]
            dist.broadcast(eng.flat.P, src=0, group=self.arena.group)
            dist.broadcast(eng.flat.S, src=0, group=self.arena.group)
            eng.sync_shadows()
            self.arena.barrier()
        if self.weights is None:
            self.weights = [1.0 / self.world] * self.world
        fl = eng.flat
        ranges, s_mask, self.beta_shared = self._share_plan(fl)
        self.aggregator = FusedAggregator(world=self.world, weights=self.weights, flat_n=fl.beta_off,
                                          beta_off=fl.beta_off, s_n=fl.s_n, K=eng.K, V=eng.V, Kp=eng.Kp,
                                          ldb=(eng.beta_bf.stride(0) if eng.is_prod else 0), lr=m.lr,
                                          betas=eng.betas, eps=eng.adam_eps, solver=eng.solver, ranges=ranges,
                                          s_mask=s_mask,
                                          beta_shared=self.beta_shared,
                                          local_gscale=1.0 / max(self.weights[self.rank], 1e-30))
        self.aggregator.add_rank(self.arena, self.rank, M=fl.M, V2=fl.V2, has_shadows=eng.is_prod,
                                 ld_gb=(eng.beta_bf.stride(0) if eng.is_prod else 0))
        self._beta_master_partial = False
        # Overlap: the beta part of the aggregation (the bulk of the bytes) runs on a few CTAs of the d-theta
        # sweep, i.e. concurrently with ~1 ms of backward compute that no longer touches beta.
        self._overlap = None
        if not hasattr(self, "_ctm_trick"):
            self._ctm_trick = False
        # CTAs that aggregate from the start of the d-theta sweep; every sweep CTA joins once it runs out of sweep work
        # (measured, large config: N=2 4 / 8 / 16 CTAs -> 3.33 / 3.30 / 3.26 ms/step; N=8 2 / 4 / 8 / 16 -> 3.25 / 3.28 /
        # 3.27 / 3.27, i.e. flat)
        nvls = bool(self.aggregator.args.ranks[0].P_mc)
        ctas = int(os.environ.get("GFEDNTM_AGG_CTAS", "16"))
        if (self.world > 1 and self.agg == "grad" and ctas >= 0 and eng.is_prod and eng.decoder.folded
                and isinstance(self.arena, SymmetricArena) and self.beta_shared and eng.V % 8 == 0
                and os.environ.get("GFEDNTM_NO_OVERLAP", "0") != "1"):
            self._overlap = self.aggregator.overlap_handle(eng.ctr, ctas)

    CTM_KEYS = ("inf_net.adapt_bert.weight", "inf_net.adapt_bert.bias", "inf_net.input_layer.weight",
                "inf_net.input_layer.bias")

    def _share_plan(self, fl):
        """(ranges over the flat region, BatchNorm statistics mask, beta shared?) from ``grads_to_share``.

        CombinedTM in gradient mode: the gradients of ``adapt_bert`` and of the contextual half of the first layer
        are linear images of the small gradient wrt the collapsed weight (engine/fused.py), so the clients sum
        THAT (``allreduce_small``, 150 KB) and every rank applies the identical Adam step itself; those ranges
        are marked *replicated* and never cross NVLink."""
        eng = self.model.engine
        keys = set(self.grads_to_share) if self.grads_to_share else None
        self._ctm_trick = (eng.kind == "combined" and self.agg == "grad" and self.world > 1
                           and (keys is None or all(k in keys for k in self.CTM_KEYS))
                           and (eng.V * eng.hidden[0]) % 4 == 0 and "AR" in self.arena.offsets
                           and os.environ.get("GFEDNTM_NO_CTM_TRICK", "0") != "1")
        if keys is None and not self._ctm_trick:
            return None, 63, True
        if keys is None:
            keys = set(fl.segments) | {"__all_stats__"}
        known = set(fl.segments) | {f"{bn}.{st}" for bn in ("inf_net.f_mu_batchnorm", "inf_net.f_sigma_batchnorm",
                                                            "beta_batchnorm")
                                    for st in ("running_mean", "running_var", "num_batches_tracked")}
        unknown = sorted(keys - known - {"__all_stats__"})
        if unknown:
            # (the reference raises KeyError in get_gradients for such keys)
            import logging

            logging.getLogger(__name__).warning("grads_to_share names unknown tensors (ignored): %s", unknown)
        ranges = []
        segs = sorted((s for n, s in fl.segments.items() if n != "beta"), key=lambda s: s.offset)
        cur = 0
        for sg in segs:
            end = (sg.offset + sg.numel + 3) // 4 * 4
            if sg.offset > cur:  # alignment padding: harmless either way
                ranges.append((cur, sg.offset - cur, True))
            kind = int(sg.name in keys)
            if self._ctm_trick and sg.name.startswith("inf_net.adapt_bert."):
                kind = 3   # replicated, updated by eng.ctm_adam_step (keeps the operand shadows in step)
            if self._ctm_trick and sg.name == "inf_net.input_layer.weight":
                half = eng.V * eng.hidden[0]  # storage is [2V, H0]: BoW half aggregated, contextual half replicated
                ranges.append((sg.offset, half, 1))
                ranges.append((sg.offset + half, half, 3))
                if end - sg.offset > 2 * half:   # label rows (supervised CTM): aggregated like the BoW half
                    ranges.append((sg.offset + 2 * half, end - sg.offset - 2 * half, 1))
            else:
                ranges.append((sg.offset, end - sg.offset, kind))
            cur = end
        if cur < fl.beta_off:
            ranges.append((cur, fl.beta_off - cur, True))
        s_mask = 0
        blocks = [("inf_net.f_mu_batchnorm", "running_mean"), ("inf_net.f_mu_batchnorm", "running_var"),
                  ("inf_net.f_sigma_batchnorm", "running_mean"), ("inf_net.f_sigma_batchnorm", "running_var"),
                  ("beta_batchnorm", "running_mean"), ("beta_batchnorm", "running_var")]
        for i, (bn, st) in enumerate(blocks):
            if f"{bn}.{st}" in keys or "__all_stats__" in keys:
                s_mask |= 1 << i
        return ranges, s_mask, ("beta" in keys)

    def replicated_grad_names(self):
        """Parameters whose gradient buffer already holds the GLOBAL gradient after the backward pass (CombinedTM
        in gradient mode; of ``inf_net.input_layer.weight`` only the contextual half, columns V..2V)."""
        if not getattr(self, "_ctm_trick", False):
            return ()
        return ("inf_net.adapt_bert.weight", "inf_net.adapt_bert.bias", "inf_net.input_layer.weight")

    def materialize(self) -> None:
        """Makes this rank's fp32 parameters complete.  The overlapped aggregation keeps the fp32 master copy
        of beta on the owner of each vocabulary range and ships only bf16 operand shadows; before anything
        reads all of beta (state_dict, topics, checkpoints) the other ranges are fetched from their owners.
        Collective: every rank calls it between steps."""
        if self.world > 1 and self._overlap is not None and self._beta_master_partial:
            self.arena.barrier()
            self.aggregator.gather_beta()
            self.arena.barrier()
            self._beta_master_partial = False

    # ------------------------------------------------------------------------------------------
    def step(self, bow: CsrBatch, x_ctx: Optional[torch.Tensor] = None, *, after_backward=None,
             labels: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One synchronous global step (local minibatch + aggregation).  Returns this client's summed
        minibatch loss (device tensor).  ``after_backward`` (verification hook): called between the local
        backward pass and the closing aggregation launch; forces the eager (non-graph) path."""
        from ..utils.observability import nvtx_range

        with nvtx_range("gfedntm.global_step"), self.timer.region("global_step"):
            loss = self._step(bow, x_ctx, after_backward, labels)
        if self.metrics.path and self.global_step % self.metrics_every == 0:
            t = self.timer.summary()  # (one synchronisation per logging interval)
            self.metrics.log(step=self.global_step, rank=self.rank, world=self.world, agg=self.agg,
                             loss_per_doc=float(loss.item()) / max(bow.B, 1), ms_per_step=t.get("global_step"),
                             docs_per_s=(1e3 * bow.B * self.world / t["global_step"]) if t.get("global_step") else None)
        return loss

    def _step(self, bow: CsrBatch, x_ctx, after_backward, labels=None) -> torch.Tensor:
        eng = self.model.engine
        self.global_step += 1
        if self.world == 1:
            return eng.train_step(bow, x_ctx, labels=labels)
        if self._overlap is not None and self.agg == "grad":
            self._beta_master_partial = True

        side = eng._side is not None   # CSC transpose of the batch on the engine's side stream
        key = ("fed", self.agg, bow.indptr.data_ptr(), bow.indices.data_ptr(), bow.B,
               x_ctx.data_ptr() if x_ctx is not None else 0, labels.data_ptr() if labels is not None else 0)
        # the step's graph also copies the loss to a pinned host slot (engine.loss_slot), behind the forward pass
        eng.loss_slot = slot = eng.loss_slot_for(key if after_backward is None else "eager")

        def body():
            if self.agg == "grad" and self._overlap is not None:
                # (inc_opt=1: epoch / Adam step t is set by the first launch of the step; the overlapped part reads it)
                eng.forward_backward(bow, x_ctx, training=True, gscale=self.weights[self.rank], fed=self._overlap,
                                     labels=labels, inc_opt=1, side_stream=side, loss_host=slot)
                if after_backward is not None:
                    after_backward()
                self._launch("grad", eng.step_count + 1, parts=1)
            elif self.agg == "grad":
                hook = None
                if self._ctm_trick:  # (the small all-reduce inside the backward pass uses epoch t)
                    hook = lambda buf: self.aggregator.allreduce_small(buf, eng.ctr)  # noqa: E731
                eng.forward_backward(bow, x_ctx, training=True, gscale=self.weights[self.rank], ctm_reduce_hook=hook,
                                     labels=labels, inc_opt=1, side_stream=side, loss_host=slot)
                if after_backward is not None:
                    after_backward()
                if not self.beta_shared:  # beta stays client-local: its own Adam on the unweighted gradient
                    eng.optimizer_step_beta(gscale=1.0 / max(self.weights[self.rank], 1e-30))
                self._launch("grad", eng.step_count + 1)
                if self._ctm_trick:
                    eng.ctm_adam_step(gscale=1.0)   # identical update on every rank from the summed small gradient
                else:
                    eng._refresh_ctm_shadows()
                eng._refresh_ctx_shadows()
            else:
                eng.forward_backward(bow, x_ctx, training=True, labels=labels, inc_opt=1, side_stream=side, loss_host=slot)
                eng.optimizer_step(ctr_done=True)
                if after_backward is not None:
                    after_backward()
                self._launch("param", eng.step_count)
                eng._refresh_ctm_shadows()
                eng._refresh_ctx_shadows()
            eng.join_side()

        if not eng.use_graphs or after_backward is not None:
            body()
            if self.agg == "grad":
                eng.step_count += 1
            return eng.loss
        eager_warmup = key not in eng._warm
        loss = eng.run_graphed(key, body, n_opt_steps=1)
        if eager_warmup and self.agg == "grad":
            eng.step_count += 1  # (agg="param": optimizer_step() advanced the host mirror itself)
        return loss

    def _launch(self, mode: str, step: int, parts: int = 3) -> None:
        eng = self.model.engine
        capturing = torch.cuda.is_current_stream_capturing()
        if not capturing:
            s = torch.cuda.Event(enable_timing=True)
            e = torch.cuda.Event(enable_timing=True)
            s.record()
        self.aggregator.launch(mode=mode, step=step, step_ctr=eng.ctr, parts=parts)
        if not capturing:
            e.record()
            self._agg_events.append((s, e))
            if len(self._agg_events) > 256:
                self._agg_events = self._agg_events[-128:]

    def last_aggregation_us(self) -> Optional[float]:
        """Mean device time of the aggregation kernel over the recorded steps (needs a prior sync)."""
        if not self._agg_events:
            return None
        torch.cuda.synchronize()
        ev = self._agg_events[-64:]
        return 1e3 * sum(s.elapsed_time(e) for s, e in ev) / len(ev)

    def describe_aggregation(self) -> Optional[dict]:
        """What the per-step aggregation of this trainer consists of (for bench / log lines)."""
        if self.world == 1 or self.aggregator is None:
            return None
        return {"mode": self.agg, "overlap_ctas": getattr(self._overlap, "ctas", 0),
                "nvls_multicast": bool(self.aggregator.args.ranks[0].P_mc)}

    def working_set_bytes(self) -> int:
        eng = self.model.engine
        n = 4 * eng.flat.total * 4
        if eng.is_prod:
            n += eng.beta_bf.numel() * 2 + eng.betaT_bf.numel() * 2
        return n
