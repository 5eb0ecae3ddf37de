"""Host-memory restatement of the fused aggregation (``csrc/fabric.cu``) over any ``torch.distributed``
backend (gloo on a GPU-less CI host).

Same decomposition as the kernel, with collectives as the transport instead of NVLink peer memory:

* ``agg="grad"``: the owner of a contiguous slice of the flat parameter space sums the clients' (already
  weighted) gradients, applies ONE Adam step with moments that exist only for its slice, and every rank
  receives every slice (reduce-scatter -> sharded Adam -> all-gather).
* ``agg="param"``: weighted mean of the locally updated parameters (reference semantics,
  src/federation/server.py:477-487).
* BatchNorm running statistics are averaged with the client weights in both modes.
* ``shared`` (optional boolean mask over the flat space = ``grads_to_share`` of the reference,
  federated_model.py:98-131): unshared elements stay client-local — their own Adam step on the client's
  unweighted gradient in ``agg="grad"``, untouched by the averaging in ``agg="param"``.

It exists so the *semantics* (slice ownership, sharded optimiser state, weighting) are testable without
GPUs (``tests/test_cpu_multiproc.py``), and as a slow fallback data plane for hosts without peer access.
"""
from __future__ import annotations

import math
from typing import Optional, Sequence, Tuple

import torch


def owner_slice(n: int, world: int, rank: int, granule: int = 4) -> Tuple[int, int]:
    """Contiguous slice of ``[0, n)`` owned by ``rank`` (same rule as the kernel: ``granule``-element vectors
    split evenly, the last owner takes the remainder)."""
    nv = (n + granule - 1) // granule
    per = (nv + world - 1) // world
    lo = min(rank * per * granule, n)
    hi = min((rank + 1) * per * granule, n)
    return lo, hi


class HostAggregator:
    def __init__(self, n_params: int, weights: Sequence[float], *, lr: float, betas=(0.9, 0.999), eps: float = 1e-8,
                 weight_decay: float = 0.0, group=None, shared: Optional[torch.Tensor] = None) -> None:
        import torch.distributed as dist

        self.group = group if group is not None else dist.group.WORLD
        self.world = dist.get_world_size(self.group)
        self.rank = dist.get_rank(self.group)
        assert len(weights) == self.world
        self.weights = [float(w) for w in weights]
        self.n = int(n_params)
        self.lo, self.hi = owner_slice(self.n, self.world, self.rank)
        self.lr, self.b1, self.b2, self.eps, self.wd = lr, betas[0], betas[1], eps, weight_decay
        # sharded optimiser state: only this rank's slice
        self.m = torch.zeros(self.hi - self.lo)
        self.v = torch.zeros(self.hi - self.lo)
        self.step = 0
        self.shared = None
        if shared is not None and not bool(shared.all()):
            self.shared = shared.to(torch.bool).clone()
            # client-local optimiser state of the unshared elements (full length, used where ~shared)
            self.m_loc = torch.zeros(self.n)
            self.v_loc = torch.zeros(self.n)

    def _adam(self, p, g, m, v):
        if self.wd != 0.0:
            g = g + self.wd * p
        m.mul_(self.b1).add_(g, alpha=1.0 - self.b1)
        v.mul_(self.b2).addcmul_(g, g, value=1.0 - self.b2)
        bc1 = 1.0 - self.b1 ** self.step
        bc2 = 1.0 - self.b2 ** self.step
        return p - (self.lr / bc1) * (m / (v.sqrt() / math.sqrt(bc2) + self.eps))

    # ------------------------------------------------------------------------------------------
    def _all_gather_slices(self, flat: torch.Tensor, mine: torch.Tensor) -> None:
        import torch.distributed as dist

        per = max(owner_slice(self.n, self.world, r)[1] - owner_slice(self.n, self.world, r)[0]
                  for r in range(self.world))
        pad = torch.zeros(per, dtype=flat.dtype)
        pad[: mine.numel()] = mine
        out = [torch.zeros(per, dtype=flat.dtype) for _ in range(self.world)]
        dist.all_gather(out, pad, group=self.group)
        for r in range(self.world):
            lo, hi = owner_slice(self.n, self.world, r)
            flat[lo:hi] = out[r][: hi - lo]

    def aggregate_grad(self, params: torch.Tensor, weighted_grad: torch.Tensor,
                       stats: Optional[torch.Tensor] = None) -> None:
        """``weighted_grad`` = w_rank * local gradient (the weight is folded into the backward pass, as in the
        engine).  ``params`` is updated in place on every rank; ``stats`` (BatchNorm running statistics of
        this client) is replaced by the weighted mean."""
        import torch.distributed as dist

        assert params.numel() == self.n and weighted_grad.numel() == self.n
        g = weighted_grad.clone()
        dist.all_reduce(g, group=self.group)  # transport of the reduce-scatter; only [lo, hi) is used below
        self.step += 1
        local_new = None
        if self.shared is not None:  # client-local step on the unweighted own gradient
            local_new = self._adam(params.clone(), weighted_grad / self.weights[self.rank], self.m_loc, self.v_loc)
        p = self._adam(params[self.lo:self.hi].clone(), g[self.lo:self.hi], self.m, self.v)
        self._all_gather_slices(params, p)
        if local_new is not None:
            params[~self.shared] = local_new[~self.shared]
        if stats is not None:
            self._average_stats(stats)

    def aggregate_param(self, params: torch.Tensor, stats: Optional[torch.Tensor] = None) -> None:
        import torch.distributed as dist

        w = params * self.weights[self.rank]
        dist.all_reduce(w, group=self.group)
        if self.shared is not None:
            params[self.shared] = w[self.shared]
        else:
            params.copy_(w)
        if stats is not None:
            self._average_stats(stats)

    def _average_stats(self, stats: torch.Tensor) -> None:
        import torch.distributed as dist

        s = stats * self.weights[self.rank]
        dist.all_reduce(s, group=self.group)
        stats.copy_(s)
