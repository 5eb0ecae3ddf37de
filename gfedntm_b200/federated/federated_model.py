"""Step-at-a-time trainer base shared by the federated AVITM / CTM wrappers.
Ref: src/models/federated/federated_model.py:17-197 (same method names and bookkeeping fields).

A federated client does one local minibatch step per global iteration (``train_mb_delta``), ships
the tensors named in ``grads_to_share`` and later installs the aggregate (``deltaUpdateFit``).  Two
payload semantics are supported:

* ``share="param"`` (reference behaviour): post-``optimizer.step()`` values of the shared state-dict
  entries; the server averages them weighted by corpus size.
* ``share="grad"``: the gradients of the shared parameters (buffers are still shipped by value); the
  server applies one global optimiser step and returns parameter values.  The local optimiser step is
  skipped for shared parameters in this mode.
"""
from __future__ import annotations

import datetime
import logging
from typing import Dict, List, Optional

import numpy as np
import torch

from ..utils.auxiliary_functions import save_model_as_npz
from ..utils.metrics import thresholded_thetas


class FederatedModel:
    def __init__(self, tm_params: dict, grads_to_share: List[str], logger=None, share: str = "param") -> None:
        self.tm_params = tm_params
        if logger is None:
            logging.basicConfig(level="INFO")
            logger = logging.getLogger("FederatedModel")
        self.logger = logger
        self.grads_to_share = list(grads_to_share)
        if share not in ("param", "grad"):
            raise ValueError("share must be 'param' or 'grad'")
        self.share = share

        self.model_dir = None
        self.train_data = None
        self.current_mb = -1
        self.current_epoch = -1
        self.samples_processed = -1
        self.train_loss = 0.0
        self.train_loader = None
        self.train_loader_iter = None
        self.current_batch_sample = None
        self.loss: Optional[torch.Tensor] = None
        self._last_batch_size = 0
        self.finished = False

        self.topics = None
        self.thetas = None
        self.betas = None

    # ------------------------------------------------------------------------------------------
    # client-side training
    # ------------------------------------------------------------------------------------------
    def preFit(self, train_data, save_dir=None) -> None:
        """Prepares step-wise training: shuffling loader, first minibatch, train mode."""
        self.model_dir = save_dir
        self.train_data = train_data
        self.current_mb = 0
        self.current_epoch = 0
        self.samples_processed = 0
        self.train_loss = 0.0
        self.train_loader = self.make_loader(train_data, shuffle=True)
        self._ensure_engine(self.train_loader)
        self.train_loader_iter = iter(self.train_loader)
        self.current_batch_sample = next(self.train_loader_iter)
        self.model.train()

    def _shared_keys(self) -> List[str]:
        sd = self.model.state_dict()
        keys = [k for k in self.grads_to_share if k in sd]
        missing = [k for k in self.grads_to_share if k not in sd]
        if missing:
            # the reference raises KeyError here (federated_model.py:112-113), which makes its shipped
            # config unusable for anything but CombinedTM; entries a model does not have are skipped
            self.logger.debug("grads_to_share entries not present in this model: %s", missing)
        return keys

    def get_gradients(self) -> dict:
        """Payload for the server: bookkeeping + the shared tensors."""
        params = {"current_mb": self.current_mb, "current_epoch": self.current_epoch,
                  "num_epochs": self.num_epochs}
        sd = self.model.state_dict()
        named = dict(self.model.named_parameters())
        for key in self._shared_keys():
            if self.share == "grad" and key in named and key in getattr(self, "_grad_payload", {}):
                params[key] = self._grad_payload[key]
            else:
                params[key] = sd[key]
        return params

    def set_gradients(self, modelStateDict: Dict[str, torch.Tensor]) -> None:
        """Installs the aggregated tensors into the local model."""
        local = self.model.state_dict()
        with torch.no_grad():
            for key in self._shared_keys():
                if key not in modelStateDict:
                    continue
                src = torch.as_tensor(modelStateDict[key]).to(local[key].device)
                local[key].copy_(src.to(local[key].dtype).reshape(local[key].shape))
        self.parameters_changed()

    def train_mb_delta(self) -> dict:
        """One local minibatch step; returns :meth:`get_gradients`."""
        mb = self.current_batch_sample
        self._last_batch_size = mb.size
        if self.share == "grad":
            self.loss = self.gradient_minibatch(mb)
        else:
            self.loss = self.train_minibatch(mb)
        return self.get_gradients()

    def gradient_minibatch(self, mb) -> torch.Tensor:
        """``share='grad'`` clients: forward + backward, gradients stay available for the payload, then a
        local optimiser step.  The step only matters for tensors that are NOT shared (they keep training
        locally); shared tensors are overwritten by the server's result in ``deltaUpdateFit``."""
        if self.backend == "fused":
            loss = self.engine.forward_backward(mb.bow, mb.x_ctx, training=True, labels=mb.labels).clone()
            self._grad_payload = {n: p.grad.detach().clone() for n, p in self.model.named_parameters()
                                  if n in self.grads_to_share}
            self.engine.optimizer_step()
            return loss
        self.model.zero_grad()
        loss = self._eager_forward_loss(mb)
        loss.backward()
        self._grad_payload = {n: p.grad.detach().clone() for n, p in self.model.named_parameters()
                              if n in self.grads_to_share and p.grad is not None}
        self.optimizer.step()
        return loss.detach()

    def _epoch_end(self, save_dir: Optional[str]) -> None:
        """Hook for per-epoch work (the CTM wrapper writes a snapshot)."""

    def deltaUpdateFit(self, modelStateDict: dict, save_dir: str, n_samples: int = 20) -> None:
        """Installs the aggregate and advances the local iterator / epoch counters.
        Ref: federated_avitm.py:85-145, federated_ctm.py:116-190."""
        self.set_gradients(modelStateDict)
        loss_val = float(self.loss.item()) if self.loss is not None else float("nan")
        self.samples_processed += self._last_batch_size
        self.train_loss += loss_val
        self.logger.info("-- -- Minibatch {} loss {} / samples processes {}".format(
            self.current_mb, loss_val, self.samples_processed))
        self.current_mb += 1
        try:
            self.current_batch_sample = next(self.train_loader_iter)
        except StopIteration:
            self._epoch_end(save_dir)
            # like the reference, the running loss is divided by the cumulative sample count and is
            # not reset between epochs (federated_avitm.py:106-138)
            self.train_loss /= max(self.samples_processed, 1)
            self.logger.info("Epoch: [{}/{}]\tSamples: [{}/{}]\tTrain Loss: {}\tTime: {}".format(
                self.current_epoch + 1, self.num_epochs, self.samples_processed,
                len(self.train_data) * self.num_epochs, self.train_loss, datetime.datetime.now()))
            if self.current_epoch == 0 or self.train_loss < self.best_loss_train:
                self.best_loss_train = min(self.best_loss_train, self.train_loss)
            self.best_components = self.model.beta
            self.train_loader_iter = iter(self.train_loader)
            self.current_batch_sample = next(self.train_loader_iter)
            self.current_epoch += 1
        if self.current_epoch >= self.num_epochs and not self.finished:
            self.finished = True
            self.logger.info("Epoch end reached")
            self.training_doc_topic_distributions = self.get_doc_topic_distribution(self.train_data, n_samples)
            self.get_results_model(save_dir)

    # ------------------------------------------------------------------------------------------
    # results
    # ------------------------------------------------------------------------------------------
    def get_results_model(self, save_dir: str) -> None:
        """Client-side results: topics (top-10 words), thetas (thresholded at 3e-3, L1-normalised),
        betas (row soft-max of the topic-word matrix) -> ``<save_dir>.npz``."""
        if self.best_components is None:
            self.best_components = self.model.beta
        self.topics = self.get_topics()
        thetas = np.asarray(self.get_doc_topic_distribution(self.train_data, thr=3e-3))
        self.thetas = thresholded_thetas(thetas, 3e-3)   # (idempotent: exact float64 normalisation of the rows)
        self.betas = self.get_topic_word_distribution()
        self.logger.info(f"-- -- Saving model at {save_dir} ")
        save_model_as_npz(save_dir, self)

    def get_topics_in_server(self, save_dir: str) -> None:
        """Server-side results: only the topic-word distribution (the server holds no corpus)."""
        self.betas = self.get_topic_word_distribution()
        self.logger.info("-- -- Saving global model...")
        save_model_as_npz(save_dir, self)

    def evaluate_synthetic_model(self, vocab_size: int, gt_thetas: np.ndarray, gt_betas: np.ndarray) -> dict:
        """TSS / DSS style scores against a synthetic ground truth.  Ref: federated_avitm.py:152-193."""
        from ..utils.auxiliary_functions import convert_topic_word_to_init_size
        from ..utils.metrics import dss, tss

        all_words = ["wd" + str(w) for w in np.arange(vocab_size + 1) if w > 0]
        betas = convert_topic_word_to_init_size(vocab_size=vocab_size, model=self, model_type="avitm",
                                                ntopics=self.n_components,
                                                id2token=self.tm_params.get("id2token", self.train_data.idx2token),
                                                all_words=all_words)
        out = {"tss": tss(betas, gt_betas)}
        if self.thetas is not None and gt_thetas is not None and len(gt_thetas) == len(self.thetas):
            out["dss"] = dss(self.thetas, gt_thetas)
        self.logger.info(f"-- -- Synthetic evaluation: {out}")
        return out
