"""In-box federation launcher: N clients, one process per GPU, aggregated by the fused NVLink kernel.

This is what ``python main.py --fabric N`` runs.  It keeps the two stages of the reference
(README.md:25-31 of the reference):

1. vocabulary consensus — every rank tokenises its own corpus, the local vocabularies are
   all-gathered on the box and merged on the host (``sorted(set union)``, like server.py:270-279);
2. synchronous per-minibatch federated training — ``FederatedTrainer`` (no gRPC, no NCCL on the step).

Results are written in the reference layout: ``<save_client><id>/model_<id>_<YYYYMMDD>.npz`` per
client and ``<save_server>/global_model_<YYYYMMDD>.npz`` (rank 0, after training every rank holds the
same aggregated parameters).
"""
from __future__ import annotations

import datetime as DT
import logging
import os
from typing import Optional

import numpy as np


def _worker(rank: int, world: int, port: int, kw: dict) -> None:
    import torch
    import torch.distributed as dist
    from sklearn.feature_extraction.text import CountVectorizer

    from ..models.data_preparation import TopicModelDataPreparation
    from ..models.datasets import BOWDataset
    from .federated_avitm import FederatedAVITM
    from .federated_ctm import FederatedCTM
    from .trainer import FederatedTrainer

    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ["MASTER_PORT"] = str(port)
    torch.cuda.set_device(rank)
    device = torch.device("cuda", rank)
    dist.init_process_group("nccl", rank=rank, world_size=world, device_id=device)
    cid = rank + 1
    logging.basicConfig(level=logging.INFO,
                        format="%(asctime)s [%(name)-14.14s] [%(levelname)-5.5s] %(message)s")
    log = logging.getLogger(f"fabric-client{cid}")
    log.setLevel(logging.INFO)

    # ---- stage 1: vocabulary consensus (on-box all-gather + host merge)
    from main import load_client_corpus  # same corpus loading rules as the gRPC client

    corpus = load_client_corpus(kw["data_type"], kw["source"], cid, kw["fos"])
    embeddings = None
    if kw["data_type"] == "synthetic":
        docs = [" ".join(d) for d in corpus]
    else:
        docs = [d if isinstance(d, str) else " ".join(d) for d in corpus["bow_text"].values.tolist()]
        if "embeddings" in corpus.columns:
            emb = corpus["embeddings"].values
            embeddings = (np.array([np.array(e.split(), dtype=np.float32) for e in emb])
                          if len(emb) and isinstance(emb[0], str) else np.stack([np.asarray(e, np.float32) for e in emb]))
    cv = CountVectorizer(input="content", lowercase=True, stop_words="english", binary=False)
    cv.fit(docs)
    from ..parallel.consensus import consensus_vocabulary

    vocab_map, _ = consensus_vocabulary(list(cv.vocabulary_.keys()), len(docs))
    gcv = CountVectorizer(vocabulary=vocab_map)
    bow = gcv.transform(docs)
    idx2token = gcv.get_feature_names_out()

    params = dict(kw["training_params"])
    params["input_size"] = len(idx2token)
    params["id2token"] = dict(enumerate(idx2token))
    params["seed"] = 1234  # identical initial weights on every rank (rank 0's are broadcast anyway)
    # subset sharing (the reference's [federation] grads_to_share): names that do not exist in this model
    # (e.g. inf_net.adapt_bert.* for avitm) are dropped with a warning by the trainer
    grads = [g.strip() for g in (kw.get("grads_to_share") or []) if g.strip()]
    if kw["model_type"] == "avitm":
        data = BOWDataset(bow, idx2token)
        model = FederatedAVITM(params, grads, log, backend="fused", device=device)
    else:
        if embeddings is None:
            raise ValueError("model_type 'ctm' needs an `embeddings` column")
        params["contextual_size"] = int(embeddings.shape[1])
        data = TopicModelDataPreparation().load(embeddings, bow, params["id2token"])
        model = FederatedCTM(params, grads, log, backend="fused", device=device)
    model.train_data = data

    # ---- stage 2: synchronous per-minibatch training
    trainer = FederatedTrainer(model, n_samples=len(data), world_size=world, rank=rank, agg=kw["agg"],
                               grads_to_share=grads or None)
    loader = model.make_loader(data, shuffle=True)
    trainer.prepare(loader)
    # every rank must take the same number of global steps: the longest client's epoch length
    steps = torch.tensor([len(loader)], device=device)
    dist.all_reduce(steps, op=dist.ReduceOp.MAX)
    steps_per_epoch = int(steps.item())
    total = min(kw["max_iters"], steps_per_epoch * int(params["num_epochs"]))
    start = 0
    ckpt_dir, ckpt_every = kw.get("checkpoint_dir"), int(kw.get("checkpoint_every") or 0)
    if kw.get("resume") and ckpt_dir:
        from ..utils.checkpoint import fabric_checkpoint_path, load_fabric_checkpoint

        if os.path.exists(fabric_checkpoint_path(ckpt_dir, rank)):
            start = int(load_fabric_checkpoint(ckpt_dir, trainer, loader)["global_step"])
            log.info("resumed from %s at global step %d", ckpt_dir, start)
    it = iter(loader)
    model.model.train()
    for step in range(start, total):
        try:
            mb = next(it)
        except StopIteration:
            it = iter(loader)
            mb = next(it)
        loss = trainer.step(mb.bow, mb.x_ctx)
        if rank == 0 and (step % 100 == 0 or step == total - 1):
            log.info("global step %d/%d  client-1 minibatch loss/doc %.3f", step + 1, total,
                     float(loss.item()) / mb.size)
        if ckpt_dir and ckpt_every > 0 and (step + 1) % ckpt_every == 0 and step + 1 < total:
            from ..utils.checkpoint import save_fabric_checkpoint

            save_fabric_checkpoint(ckpt_dir, trainer, loader, global_step=step + 1)
    torch.cuda.synchronize()
    dist.barrier()
    trainer.materialize()  # complete fp32 parameters on every rank (the master copy of beta is sharded by owner)

    # ---- results (reference layout)
    model.best_components = model.model.beta
    stamp = DT.datetime.now().strftime("%Y%m%d")
    cdir = f"{kw['save_client']}{cid}"
    os.makedirs(cdir, exist_ok=True)
    model.get_results_model(os.path.join(cdir, f"model_{cid}_{stamp}"))
    if rank == 0:
        os.makedirs(kw["save_server"], exist_ok=True)
        model.thetas = None
        model.topics = None
        model.get_topics_in_server(os.path.join(kw["save_server"], f"global_model_{stamp}"))
    dist.barrier()
    dist.destroy_process_group()


def supervise(launch, *, max_restarts: int = 0, can_resume: bool = False, log=None) -> int:
    """Failure recovery of the fabric path: ``launch(resume, attempt)`` runs all ranks to completion and raises
    when one of them dies (a crashed process, or the device-side trap the other ranks take 120 s after a peer's flag
    stopped advancing, ``csrc/fabric_dev.cuh``).  The federation is synchronous -- every step needs every rank -- so
    recovery is a restart of ALL ranks from the last common checkpoint: up to ``max_restarts`` times when the run
    writes checkpoints.  Returns the number of restarts that were needed.  (The reference has no recovery: a client
    that dies leaves the server waiting)"""
    import logging

    log = log or logging.getLogger("gfedntm.fabric")
    attempt = 0
    while True:
        try:
            launch(attempt > 0, attempt)
            return attempt
        except KeyboardInterrupt:
            raise
        except BaseException as ex:  # mp.spawn raises ProcessRaisedException / ProcessExitedException
            if attempt >= max_restarts or not can_resume:
                raise
            attempt += 1
            log.warning("fabric federation failed (%s: %s); restart %d of %d from the last checkpoint",
                        type(ex).__name__, str(ex).splitlines()[0] if str(ex) else "", attempt, max_restarts)


def run_fabric_federation(*, n_clients: int, source: str, data_type: str, fos: str, model_type: str,
                          training_params: dict, max_iters: int, save_client: str, save_server: str,
                          logs_client: str, agg: str = "grad", port: Optional[int] = None,
                          grads_to_share: Optional[list] = None, checkpoint_dir: Optional[str] = None,
                          checkpoint_every: int = 0, resume: bool = False, max_restarts: int = 0) -> None:
    import torch
    import torch.multiprocessing as mp

    if torch.cuda.device_count() < n_clients:
        raise RuntimeError(f"--fabric {n_clients} needs {n_clients} GPUs, found {torch.cuda.device_count()}")
    port = port or (29400 + os.getpid() % 500)

    def launch(resume_now: bool, attempt: int) -> None:
        kw = dict(source=source, data_type=data_type, fos=fos, model_type=model_type,
                  training_params=training_params, max_iters=max_iters, save_client=save_client,
                  save_server=save_server, logs_client=logs_client, agg=agg, grads_to_share=grads_to_share,
                  checkpoint_dir=checkpoint_dir, checkpoint_every=checkpoint_every, resume=resume or resume_now)
        # (a fresh rendezvous port per attempt: the previous store may still be in TIME_WAIT)
        mp.spawn(_worker, args=(n_clients, port + attempt, kw), nprocs=n_clients, join=True)

    supervise(launch, max_restarts=max_restarts, can_resume=bool(checkpoint_dir) and checkpoint_every > 0)
