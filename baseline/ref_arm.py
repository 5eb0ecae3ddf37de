"""Reference arm of bench.py: runs the UNMODIFIED gFedNTM sources from ``baseline/_ref`` (populated by
``baseline/install_ref.sh``) through their own classes for the same metric and configuration.

Nothing of gfedntm_b200 is on this path: models, loss, optimiser, DataLoader and (for N > 1) the
step-wise federated wrappers are the reference's.  Three environment shims are applied before import,
none of which touches a reference file: ``numpy.Inf`` (removed in NumPy 2),
and import stubs for ``matplotlib`` and ``waiting`` (not installed on the box, unused on this path).

N = 1  -> ``AVITM._train_epoch`` / ``CTM._train_epoch`` (the reference's stock training path).
N > 1  -> every rank runs ``FederatedAVITM.train_mb_delta()`` / ``FederatedCTM.train_mb_delta()``
          (stock local step) and ``set_gradients`` (stock installation of the aggregate); the hop in
          between is a torch+NCCL weighted all-reduce of the shared tensors — BASELINE.md variant (c).
          The reference's own gRPC hop cannot be used for avitm (server.py:294 raises AttributeError)
          and sleeps 3 s per client per step by construction; it is reported separately in BASELINE.md.
"""
from __future__ import annotations

import logging
import os
import sys
import time
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_ROOT = os.path.join(HERE, "_ref")


def _shim_env() -> None:
    if not hasattr(np, "Inf"):
        np.Inf = np.inf  # noqa: NPY201
    for name in ("matplotlib", "matplotlib.pyplot", "waiting", "wordcloud"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                mod = types.ModuleType(name)
                if name == "waiting":
                    mod.wait = lambda *a, **k: True
                sys.modules[name] = mod
    if "matplotlib" in sys.modules and not hasattr(sys.modules["matplotlib"], "pyplot") \
            and "matplotlib.pyplot" in sys.modules:
        sys.modules["matplotlib"].pyplot = sys.modules["matplotlib.pyplot"]


def ensure_installed(source: str | None = None) -> str:
    """Make sure ``baseline/_ref`` holds the unmodified reference sources and return its path.

    ``__graft_entry__.build()`` calls this on the build box (where ``/root/reference`` is mounted) so that
    the copy travels with the tree to the GPU boxes, where the mount does not exist.  The source may be
    overridden with ``GFEDNTM_REFERENCE_SRC``.  Nothing is patched: it is ``install_ref.sh`` on demand."""
    if os.path.isdir(os.path.join(REF_ROOT, "src", "models")):
        return REF_ROOT
    source = source or os.environ.get("GFEDNTM_REFERENCE_SRC", "/root/reference")
    if not os.path.isdir(os.path.join(source, "src")):
        raise ImportError(f"baseline/_ref is empty and no reference checkout at {source} "
                          "(run baseline/install_ref.sh <path>)")
    import subprocess

    subprocess.run(["bash", os.path.join(HERE, "install_ref.sh"), source], check=True,
                   stdout=subprocess.DEVNULL, stderr=subprocess.DEVNULL)
    if not os.path.isdir(os.path.join(REF_ROOT, "src", "models")):
        raise ImportError("baseline/install_ref.sh did not produce baseline/_ref/src")
    return REF_ROOT


def _import_reference():
    ensure_installed()
    _shim_env()
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    from src.models.base.contextualized_topic_models.ctm_network.ctm import CTM as RefCTM
    from src.models.base.contextualized_topic_models.datasets.dataset import CTMDataset as RefCTMDataset
    from src.models.base.pytorchavitm.avitm_network.avitm import AVITM as RefAVITM
    from src.models.base.pytorchavitm.datasets.bow_dataset import BOWDataset as RefBOWDataset
    from src.models.federated.federated_avitm import FederatedAVITM as RefFedAVITM
    from src.models.federated.federated_ctm import FederatedCTM as RefFedCTM

    return RefAVITM, RefCTM, RefBOWDataset, RefCTMDataset, RefFedAVITM, RefFedCTM


def run_reference_arm(args, cfg, ClockSampler, dist_env, make_corpus, allreduce_max) -> dict:
    import torch
    import torch.distributed as dist
    from torch.utils.data import DataLoader

    try:
        RefAVITM, RefCTM, RefBOW, RefCTMData, RefFedAVITM, RefFedCTM = _import_reference()
    except Exception as ex:  # pragma: no cover
        return {"impl": "reference", "unavailable": f"{type(ex).__name__}: {ex}"[:200]}

    rank, world, local = dist_env()
    torch.cuda.set_device(local)
    device = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=device)
    B, V, K = cfg["B"], cfg["V"], cfg["K"]
    n_batches = 2 if V * B > 2e8 else 4
    counts, ctx = make_corpus(cfg, B * n_batches, seed=1234 + rank, device=f"cuda:{local}")
    dense = np.asarray(counts.todense(), dtype=np.float32)  # the reference's input format: dense [n_docs, V]
    logger = logging.getLogger("reference")
    logger.setLevel(logging.WARNING)
    is_ctm = cfg["family"] == "ctm"
    idx2token = [f"wd{i}" for i in range(V)]
    common = dict(input_size=V, n_components=K, model_type=cfg["model_type"], hidden_sizes=tuple(cfg["hidden"]),
                  activation="softplus", dropout=0.2, learn_priors=True, batch_size=B, lr=2e-3, momentum=0.99,
                  solver="adam", num_epochs=1, reduce_on_plateau=False, topic_prior_mean=0.0,
                  topic_prior_variance=None, num_samples=20, num_data_loader_workers=0, verbose=False)
    note = None
    if world == 1:
        if is_ctm:
            model = RefCTM(logger=logger, contextual_size=cfg["C"], inference_type=cfg["inference_type"], **common)
            data = RefCTMData(ctx, dense, idx2token, None)  # (X_contextual, X_bow, idx2token, qt)
        else:
            model = RefAVITM(logger=logger, **common)
            data = RefBOW(dense, idx2token)
        fed = None
    else:
        tm = dict(common)
        tm["contextual_size"] = cfg.get("C", 0)
        shared = None  # all state-dict entries
        if is_ctm:
            if cfg["inference_type"] != "combined":
                return {"impl": "reference",
                        "unavailable": "the reference's federated CTM is always CombinedTM (federated_ctm.py:24-45)"}
            model = RefFedCTM(tm, [], logger)
            data = RefCTMData(ctx, dense, idx2token, None)  # (X_contextual, X_bow, idx2token, qt)
        else:
            model = RefFedAVITM(tm, [], logger)
            data = RefBOW(dense, idx2token)
        model.grads_to_share = list(model.model.state_dict().keys())
        fed = model
        note = ("reference model code + torch NCCL weighted all-reduce in place of its gRPC hop "
                "(server.py:294 crashes for avitm; the gRPC loop sleeps 3 s per client per step)")
    model.model.train()

    def to_batch(i0):
        sl = slice(i0 * B, (i0 + 1) * B)
        if is_ctm:
            return {"X_bow": torch.from_numpy(dense[sl]).to(device), "X_contextual": torch.from_numpy(ctx[sl]).to(device)}
        return {"X": torch.from_numpy(dense[sl]).to(device)}

    staged = [to_batch(i) for i in range(n_batches)]
    weights = None
    if world > 1:
        n = torch.tensor([float(len(data))], device=device)
        alln = [torch.zeros_like(n) for _ in range(world)]
        dist.all_gather(alln, n)
        tot = sum(float(x.item()) for x in alln)
        weights = [float(x.item()) / tot for x in alln]

    def fed_step(batch):
        """Stock local step + NCCL weighted average + stock installation of the aggregate."""
        fed.current_batch_sample = batch
        upd = fed.train_mb_delta()
        agg = {}
        for k in fed.grads_to_share:
            t = upd[k].detach().to(torch.float32) * weights[rank]
            dist.all_reduce(t)
            agg[k] = t.to(upd[k].dtype)
        fed.set_gradients(agg)

    def run_steps(batches):
        if fed is None:
            model._train_epoch(batches)  # stock epoch loop (forward, _loss, backward, step, loss.item())
        else:
            for b in batches:
                fed_step(b)

    def barrier():
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    run_steps([staged[i % n_batches] for i in range(args.warmup)])
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    run_steps([staged[i % n_batches] for i in range(args.steps)])
    e.record()
    barrier()
    ms = allreduce_max(s.elapsed_time(e), world, device)

    # ---- end to end: the reference's own Dataset + torch DataLoader from host memory, loss read back
    loader = DataLoader(data, batch_size=B, shuffle=True, num_workers=0, drop_last=True)
    e2e_steps = max(1, min(args.steps, len(loader)))
    it = iter(loader)
    first = [next(it)]
    run_steps(first)
    barrier()
    t0 = time.perf_counter()
    done = 0
    it = iter(loader)
    while done < e2e_steps:
        try:
            batch = next(it)
        except StopIteration:
            it = iter(loader)
            batch = next(it)
        run_steps([batch])
        done += 1
    barrier()
    e2e_s = allreduce_max(time.perf_counter() - t0, world, device)
    clocks = sampler.stop() if rank == 0 else None
    docs = world * B * args.steps
    h2d = B * V * 4 + (B * cfg["C"] * 4 if is_ctm else 0)
    out = {
        "metric": "ProdLDA/CTM docs/sec (federated, one client per GPU)", "value": docs / (ms / 1e3),
        "unit": "docs/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "fp32", "data": "synthetic LDA corpus (same generator as the `ours` arm), random-init weights",
        "config": {"model": cfg["desc"], "name": args.config, "global_batch": world * B, "seq_len": None,
                   "vocab": V, "topics": K, "hidden": list(cfg["hidden"]), "parallelism": f"fed{world}",
                   "l2": "dense [B,V] fp32 activations larger than L2" if B * V * 4 > 2.6e8 else "no L2 flush"},
        "e2e": {"value": world * B * e2e_steps / e2e_s, "unit": "docs/s", "h2d_bytes_per_step": int(h2d),
                "d2h_bytes_per_step": 4, "ms_per_step": 1e3 * e2e_s / e2e_steps, "steps": e2e_steps},
        "gpu_launches": None, "clocks": clocks, "impl": "reference",
    }
    if note:
        out["variant"] = note
    if world > 1:
        dist.destroy_process_group()
    return out
