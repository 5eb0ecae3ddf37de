"""Mid-training checkpoint / resume (the reference only saves final results).

A checkpoint holds everything a client (or a rank of the in-box federation) needs to continue exactly
where it stopped: model state dict, optimiser state (Adam moments + step), the trainer's bookkeeping
(epoch / minibatch / samples / losses), the loader's RNG state and the Philox seed/counters of the
fused engine.  Files are written atomically (tmp + rename) so a crash during save never corrupts the
previous checkpoint.
"""
from __future__ import annotations

import os
import tempfile
from typing import Optional

import torch


def save_training_checkpoint(path: str, trainer, loader=None, extra: Optional[dict] = None) -> None:
    sd = {k: v.detach().cpu().contiguous() for k, v in trainer.model.state_dict().items()}
    opt = trainer.optimizer.state_dict()
    opt_cpu = {"param_groups": opt["param_groups"],
               "state": {k: {n: (t.detach().cpu() if isinstance(t, torch.Tensor) else t) for n, t in st.items()}
                         for k, st in opt["state"].items()}}
    book = {k: getattr(trainer, k) for k in ("current_mb", "current_epoch", "samples_processed", "train_loss",
                                             "best_loss_train", "nn_epoch", "finished") if hasattr(trainer, k)}
    eng = getattr(trainer, "engine", None)
    engine_state = None
    if eng is not None:
        engine_state = {"seed": eng.seed, "rng_calls": eng.rng_calls, "step_count": eng.step_count,
                        "ctr": eng.ctr.detach().cpu().tolist()}
    payload = {"format": "gfedntm_b200.ckpt.v1", "state_dict": sd, "optimizer": opt_cpu, "bookkeeping": book,
               "engine": engine_state, "torch_rng": torch.get_rng_state(),
               "loader_rng": loader.rng.bit_generator.state if loader is not None else None,
               "extra": extra or {}}
    d = os.path.dirname(os.path.abspath(path))
    os.makedirs(d, exist_ok=True)
    fd, tmp = tempfile.mkstemp(dir=d, suffix=".tmp")
    with os.fdopen(fd, "wb") as f:
        torch.save(payload, f)
    os.replace(tmp, path)


def load_training_checkpoint(path: str, trainer, loader=None) -> dict:
    with open(path, "rb") as f:
        payload = torch.load(f, map_location="cpu", weights_only=False)
    if payload.get("format") != "gfedntm_b200.ckpt.v1":
        raise ValueError("not a gfedntm_b200 training checkpoint")
    trainer.model.load_state_dict(payload["state_dict"])
    trainer.parameters_changed()
    if payload["optimizer"].get("state"):
        trainer.optimizer.load_state_dict(payload["optimizer"])
    for k, v in payload["bookkeeping"].items():
        setattr(trainer, k, v)
    eng = getattr(trainer, "engine", None)
    if eng is not None and payload.get("engine"):
        es = payload["engine"]
        eng.seed, eng.rng_calls = es["seed"], es["rng_calls"]
        eng.set_step_count(es["step_count"])
        eng.ctr[0] = int(es["ctr"][0])
    torch.set_rng_state(payload["torch_rng"])
    if loader is not None and payload.get("loader_rng") is not None:
        loader.rng.bit_generator.state = payload["loader_rng"]
    return payload.get("extra", {})


# ----------------------------------------------------------------------------------------------------------------
# In-box federation (one rank per GPU, fused NVLink aggregation): one file per rank
# ----------------------------------------------------------------------------------------------------------------
def fabric_checkpoint_path(directory: str, rank: int) -> str:
    return os.path.join(directory, f"fabric_rank{rank}.ckpt")


def save_fabric_checkpoint(directory: str, trainer, loader=None, global_step: int = 0, extra: Optional[dict] = None) -> None:
    """Checkpoint of one rank of ``FederatedTrainer``.  Collective: every rank calls it at the same global step.

    Besides the model state it stores this rank's flat Adam moments as they are: in gradient mode the optimiser
    state is sharded (a rank holds the moments of the parameter slices it owns), so every rank has to restore its
    own file.  ``trainer.materialize()`` first completes the fp32 parameters (the overlapped aggregation keeps the
    master copy of beta on its owners)."""
    trainer.materialize()
    m = trainer.model
    eng = m.engine
    torch.cuda.synchronize()
    payload = {"format": "gfedntm_b200.fabric_ckpt.v1", "rank": trainer.rank, "world": trainer.world, "agg": trainer.agg,
               "global_step": int(global_step),
               "state_dict": {k: v.detach().cpu().contiguous() for k, v in m.model.state_dict().items()},
               "adam_m": eng.flat.M.detach().cpu(), "adam_v": eng.flat.V2.detach().cpu(),
               "engine": {"seed": eng.seed, "rng_calls": eng.rng_calls, "step_count": eng.step_count,
                          "ctr": eng.ctr.detach().cpu().tolist()},
               "loader_rng": loader.rng.bit_generator.state if loader is not None else None,
               "weights": list(trainer.weights), "extra": extra or {}}
    os.makedirs(directory, exist_ok=True)
    path = fabric_checkpoint_path(directory, trainer.rank)
    fd, tmp = tempfile.mkstemp(dir=directory, suffix=".tmp")
    with os.fdopen(fd, "wb") as f:
        torch.save(payload, f)
    os.replace(tmp, path)


def load_fabric_checkpoint(directory: str, trainer, loader=None) -> dict:
    """Restores one rank (after ``trainer.prepare``); returns ``{"global_step": ..., "extra": ...}``."""
    path = fabric_checkpoint_path(directory, trainer.rank)
    with open(path, "rb") as f:
        payload = torch.load(f, map_location="cpu", weights_only=False)
    if payload.get("format") != "gfedntm_b200.fabric_ckpt.v1":
        raise ValueError("not a gfedntm_b200 fabric checkpoint")
    if payload["world"] != trainer.world or payload["rank"] != trainer.rank:
        raise ValueError(f"checkpoint of rank {payload['rank']}/{payload['world']} loaded into rank "
                         f"{trainer.rank}/{trainer.world}: the sharded optimiser state would not line up")
    m = trainer.model
    eng = m.engine
    m.model.load_state_dict(payload["state_dict"])
    eng.flat.M.copy_(payload["adam_m"].to(eng.device))
    eng.flat.V2.copy_(payload["adam_v"].to(eng.device))
    es = payload["engine"]
    eng.seed, eng.rng_calls = es["seed"], es["rng_calls"]
    eng.set_step_count(es["step_count"])
    eng.ctr[0] = int(es["ctr"][0])
    eng.sync_shadows()
    if loader is not None and payload.get("loader_rng") is not None:
        loader.rng.bit_generator.state = payload["loader_rng"]
    torch.cuda.synchronize()
    trainer.arena.barrier()
    return {"global_step": payload["global_step"], "extra": payload.get("extra", {})}
