"""Mid-training checkpoint / resume (the reference only saves final results).

A checkpoint holds everything a client (or a rank of the in-box federation) needs to continue exactly
where it stopped: model state dict, optimiser state (Adam moments + step), the trainer's bookkeeping
(epoch / minibatch / samples / losses), the loader's RNG state and the Philox seed/counters of the
fused engine.  Files are written atomically (tmp + rename) so a crash during save never corrupts the
previous checkpoint.
"""
from __future__ import annotations

import os
import pickle
import tempfile
from typing import Optional

import numpy as np
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
