"""Cross-rank verification of the fused NVLink aggregation, usable inside any multi-GPU run.

Two checks (both collective: every rank must call them at the same point):

* ``ranks_identical`` — after a global step every rank must hold bit-identical replicated state
  (flat parameters, BatchNorm statistics, bf16 operand shadows): an integer checksum per buffer is
  reduced with MIN and MAX over the ranks and the two must agree.
* ``oracle_max_rel_err`` — a few extra global steps are taken while an oracle follows along: the ranks'
  weighted gradients are summed with a plain NCCL all-reduce and applied by ``torch.optim.Adam`` to a
  private copy of the parameters (``agg="grad"``), or the post-step parameters are averaged with the
  client weights (``agg="param"``, the formula of the reference's server.py:477-487).  Reported is the
  largest ``max|fused - oracle| / max|oracle|`` over all tensors and ranks.

``bench.py`` carries both in its JSON line at N > 1; ``tests/mp_fed_check.py`` asserts them.
"""
from __future__ import annotations

from typing import Dict

import torch


def _checksum(t: torch.Tensor) -> torch.Tensor:
    """Order-independent integer checksum of the raw bits (int64 sum of the 32-bit words)."""
    raw = t.detach().contiguous().view(torch.uint8)
    n = raw.numel() // 4 * 4
    return raw[:n].view(torch.int32).to(torch.int64).sum().reshape(1)


def ranks_identical(trainer) -> bool:
    import torch.distributed as dist

    eng = trainer.model.engine
    trainer.materialize()  # (overlapped aggregation: the fp32 master copy of beta lives on its owners)
    bufs = [eng.flat.P, eng.flat.S]
    if eng.is_prod:
        bufs += [eng.beta_bf, eng.betaT_bf]
    sums = torch.cat([_checksum(b) for b in bufs])
    lo, hi = sums.clone(), sums.clone()
    group = getattr(trainer.arena, "group", None)
    dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
    dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
    return bool(torch.equal(lo, hi))


def verify_against_oracle(trainer, loader, steps: int = 2) -> Dict[str, object]:
    """Runs ``steps`` extra global steps next to the NCCL + torch oracle.  Returns
    ``{"ranks_identical": bool, "oracle_max_rel_err": float, "oracle_steps": int}``."""
    import torch.distributed as dist

    model, eng = trainer.model, trainer.model.engine
    group = getattr(trainer.arena, "group", None)
    rank, w = trainer.rank, trainer.weights
    torch.cuda.synchronize()
    identical = ranks_identical(trainer)
    names = [n for n, _ in model.model.named_parameters()]
    ref = {n: eng.flat.view(eng.flat.P, n).detach().clone().contiguous().requires_grad_(True) for n in names}
    opt = torch.optim.Adam([ref[n] for n in names], lr=model.lr, betas=(model.momentum, 0.99))
    # Both optimisers continue from step t with zeroed moments (the fused moments are sharded by owner in
    # grad mode; zeroing them on every rank and in the oracle gives the same well-defined starting state
    # without reassembling the shards).  The step counter is kept: the flag epochs are derived from it.
    eng.flat.M.zero_()
    eng.flat.V2.zero_()
    for n in names:
        opt.state[ref[n]] = {"step": torch.tensor(float(eng.step_count)), "exp_avg": torch.zeros_like(ref[n]),
                             "exp_avg_sq": torch.zeros_like(ref[n])}
    worst = 0.0
    it = iter(loader)
    for _ in range(steps):
        try:
            mb = next(it)
        except StopIteration:
            it = iter(loader)
            mb = next(it)
        grabbed: Dict[str, torch.Tensor] = {}

        def grab():
            if trainer.agg == "grad":
                for n in names:
                    if n == "beta" and trainer._overlap is not None:
                        # exchange format of the overlapped path: bf16 (the oracle sums the same values)
                        grabbed[n] = trainer._overlap.gb[:, :eng.V].float().contiguous()
                    else:
                        grabbed[n] = eng.flat.view(eng.flat.G, n).detach().clone().contiguous()
            else:
                for n in names:
                    grabbed[n] = eng.flat.view(eng.flat.P, n).detach().clone().contiguous() * w[rank]
            grabbed["__S"] = eng.flat.S.detach().clone() * w[rank]

        trainer.step(mb.bow, mb.x_ctx, after_backward=grab)
        repl = trainer.replicated_grad_names() if trainer.agg == "grad" else ()
        for n in names:
            if n in repl and n != "inf_net.input_layer.weight":
                continue  # already the global gradient on every rank (summed through the collapsed weight)
            if n in repl:  # first layer [H0, 2V]: BoW half is local, contextual half already global
                loc = grabbed[n].clone()
                dist.all_reduce(loc, group=group)
                grabbed[n][:, :eng.V] = loc[:, :eng.V]
                continue
            dist.all_reduce(grabbed[n], group=group)
        dist.all_reduce(grabbed["__S"], group=group)
        if trainer.agg == "grad":
            for n in names:
                ref[n].grad = grabbed[n]
            opt.step()
        else:
            with torch.no_grad():
                for n in names:
                    ref[n].copy_(grabbed[n])
        torch.cuda.synchronize()
        trainer.materialize()
        for n in names:
            got, want = eng.flat.view(eng.flat.P, n).detach(), ref[n].detach()
            err = float((got - want).abs().max()) / max(float(want.abs().max()), 1e-30)
            worst = max(worst, err)
        s_err = float((eng.flat.S - grabbed["__S"]).abs().max()) / max(float(grabbed["__S"].abs().max()), 1e-30)
        worst = max(worst, s_err)
        identical = identical and ranks_identical(trainer)
    if hasattr(it, "close"):
        it.close()
    t = torch.tensor([worst], dtype=torch.float64, device=eng.device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return {"ranks_identical": bool(identical), "oracle_max_rel_err": float(t.item()), "oracle_steps": steps}
