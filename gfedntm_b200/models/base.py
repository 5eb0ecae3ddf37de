"""Shared trainer for the VAE topic models (AVITM and CTM families).

Two execution back ends behind one API:

* ``fused`` — ``gfedntm_b200.engine.FusedEngine``: the whole minibatch step as sm_100a kernels
  (default whenever a CUDA device is present and the configuration is supported);
* ``eager`` — plain PyTorch autograd on the ``DecoderNetwork`` module (CPU, or unsupported options
  such as a non-Adam solver or a supervised label head).

Ref: src/models/base/pytorchavitm/avitm_network/avitm.py (AVITM) and
src/models/base/contextualized_topic_models/ctm_network/ctm.py (CTM); method names, defaults and
file formats follow those classes.
"""
from __future__ import annotations

import datetime
import logging
import os
from typing import List, Optional

import numpy as np
import torch
from scipy.special import softmax
from torch import optim
from torch.optim.lr_scheduler import ReduceLROnPlateau

from ..utils.loader import CsrMinibatchLoader, Minibatch
from .early_stopping import EarlyStopping
from .networks import ACTIVATION_NAMES, DecoderNetwork, elbo_terms

SOLVERS = ("adagrad", "adam", "sgd", "adadelta", "rmsprop")


class PlateauLR:
    """``ReduceLROnPlateau(mode='min', factor=0.1, patience, threshold=1e-4 rel)`` for the fused engine, whose learning
    rate is a kernel argument rather than a ``torch.optim`` param group.  Ref: avitm.py:156-157, 418-419."""

    def __init__(self, trainer, patience: int = 10, factor: float = 0.1, threshold: float = 1e-4, min_lr: float = 0.0):
        self.trainer, self.patience, self.factor, self.threshold, self.min_lr = trainer, patience, factor, threshold, min_lr
        self.best, self.num_bad = float("inf"), 0

    def step(self, metric: float) -> None:
        metric = float(metric)
        if metric < self.best * (1.0 - self.threshold):
            self.best, self.num_bad = metric, 0
        else:
            self.num_bad += 1
        if self.num_bad > self.patience:
            new_lr = max(self.trainer.lr * self.factor, self.min_lr)
            if self.trainer.lr - new_lr > 1e-8:
                self.trainer.lr = new_lr
                if self.trainer.engine is not None:
                    self.trainer.engine.set_lr(new_lr)
                self.trainer.logger.info("reducing learning rate to %g", new_lr)
            self.num_bad = 0


class FusedAdamHandle:
    """``torch.optim.Adam``-shaped view (``state_dict`` / ``load_state_dict`` / ``param_groups``) of
    the fused engine's flat Adam state, so code that ships or restores optimiser state (the gRPC
    ``OptUpdate`` message, checkpoints) works unchanged with the fused back end."""

    def __init__(self, trainer: "NeuralTopicModel") -> None:
        self._t = trainer

    @property
    def param_groups(self):
        t = self._t
        return [{"lr": t.lr, "betas": (t.momentum, 0.99), "eps": 1e-8, "weight_decay": 0.0,
                 "amsgrad": False, "params": list(range(len(list(t.model.parameters()))))}]

    def state_dict(self) -> dict:
        t = self._t
        state = {}
        eng = t.engine
        if eng is not None and eng.step_count > 0:
            for i, (name, _) in enumerate(t.model.named_parameters()):
                state[i] = {"step": torch.tensor(float(eng.step_count)),
                            "exp_avg": eng.flat.view(eng.flat.M, name).detach().clone(),
                            "exp_avg_sq": eng.flat.view(eng.flat.V2, name).detach().clone()}
        return {"state": state, "param_groups": self.param_groups}

    def load_state_dict(self, sd: dict) -> None:
        t = self._t
        groups = sd.get("param_groups") or []
        if groups:
            g = groups[0]
            t.lr = float(g.get("lr", t.lr))
            betas = g.get("betas", (t.momentum, 0.99))
            t.momentum = float(betas[0])
            if t.engine is not None:
                t.engine.lr = t.lr
                t.engine.betas = (float(betas[0]), float(betas[1]))
        t._pending_opt_state = sd.get("state") or {}
        t._apply_pending_opt_state()

    def zero_grad(self, set_to_none: bool = False) -> None:
        if self._t.engine is not None:
            self._t.engine.flat.G.zero_()

    def step(self) -> None:
        if self._t.engine is None:
            raise RuntimeError("fused optimiser has no engine yet (call fit/preFit first)")
        self._t.engine.optimizer_step()


class NeuralTopicModel:
    """Common implementation; see :class:`AVITM` and :class:`CTM` for the public constructors."""

    _family = "avitm"

    def _setup(self, *, logger, input_size, n_components, model_type, hidden_sizes, activation, dropout,
               learn_priors, batch_size, lr, momentum, solver, num_epochs, reduce_on_plateau,
               topic_prior_mean, topic_prior_variance, num_samples, num_data_loader_workers, verbose,
               infnet="bow", contextual_size=0, label_size=0, loss_weights=None, backend="auto",
               device=None, seed=None) -> None:
        assert isinstance(input_size, int) and input_size > 0, "input_size must by type int > 0."
        assert isinstance(n_components, int) and n_components > 0, "n_components must by type int > 0."
        if self._family == "ctm":
            assert model_type in ("LDA", "prodLDA"), "model must be 'LDA' or 'prodLDA'."
        else:
            assert model_type.lower() in ("lda", "prodlda"), "model must be 'LDA' or 'prodLDA'."
        assert isinstance(hidden_sizes, tuple), "hidden_sizes must be type tuple."
        assert activation in ACTIVATION_NAMES + ("swish",), \
            "activation must be 'softplus', 'relu', 'sigmoid', 'swish', 'leakyrelu', 'rrelu', 'elu', 'selu' or 'tanh'."
        assert dropout >= 0, "dropout must be >= 0."
        assert isinstance(learn_priors, bool), "learn_priors must be boolean."
        assert isinstance(batch_size, int) and batch_size > 0, "batch_size must be int > 0."
        assert lr > 0, "lr must be > 0."
        assert isinstance(momentum, float) and 0 < momentum <= 1, "momentum must be 0 < float <= 1."
        assert solver in SOLVERS, "solver must be 'adam', 'adadelta', 'sgd', 'rmsprop' or 'adagrad'"
        assert isinstance(reduce_on_plateau, bool), "reduce_on_plateau must be type bool."
        assert isinstance(topic_prior_mean, float), "topic_prior_mean must be type float"

        if logger is None:
            logging.basicConfig(level="INFO")
            logger = logging.getLogger(type(self).__name__)
        self.logger = logger
        self.input_size = input_size
        self.n_components = n_components
        self.model_type = model_type
        self.hidden_sizes = hidden_sizes
        self.activation = activation
        self.dropout = dropout
        self.learn_priors = learn_priors
        self.batch_size = batch_size
        self.lr = lr
        self.momentum = momentum
        self.solver = solver
        self.num_epochs = num_epochs
        self.reduce_on_plateau = reduce_on_plateau
        self.topic_prior_mean = topic_prior_mean
        self.topic_prior_variance = topic_prior_variance
        self.num_data_loader_workers = num_data_loader_workers
        self.num_samples = num_samples
        self.verbose = verbose
        self.contextual_size = contextual_size
        self.label_size = label_size
        self.infnet = infnet
        self.weights = loss_weights if loss_weights else {"beta": 1}
        self.seed = int(seed) if seed is not None else int(torch.initial_seed() % (1 << 31))

        self.best_loss_train = float("inf")
        self.model_dir = None
        self.train_data = None
        self.nn_epoch = None
        self.best_components = None
        self.validation_data = None
        self.training_doc_topic_distributions = None

        self.USE_CUDA = torch.cuda.is_available() if device is None else torch.device(device).type == "cuda"
        self.device = torch.device(device) if device is not None else torch.device(
            "cuda" if self.USE_CUDA else "cpu")

        # The reference never forwards topic_prior_mean/variance into its network (avitm.py:132-134):
        # priors are always 0 and 1 - 1/K.  Kept.
        self.model = DecoderNetwork(input_size, n_components, model_type, hidden_sizes, activation, dropout,
                                    learn_priors, infnet=infnet, contextual_size=contextual_size,
                                    label_size=label_size)
        self.model = self.model.to(self.device)
        self.early_stopping = EarlyStopping(patience=5, verbose=False)

        fused_ok = (self.device.type == "cuda" and label_size <= 256
                    and activation != "swish" and n_components <= 236 and max(hidden_sizes) <= 256
                    and len(hidden_sizes) <= 8)
        if backend == "auto":
            backend = "fused" if fused_ok else "eager"
            if not fused_ok and self.device.type == "cuda":
                why = [msg for cond, msg in (
                    (label_size > 256, "label_size>256"),
                    (activation == "swish", "activation='swish'"), (n_components > 236, "n_components>236"),
                    (max(hidden_sizes) > 256, "hidden size>256"), (len(hidden_sizes) > 8, "more than 8 hidden layers"))
                       if cond]
                self.logger.warning("this configuration (%s) is outside the fused sm_100a engine and runs on the eager "
                                    "PyTorch backend, typically 10-70x slower per step", ", ".join(why))
        if backend == "fused" and not fused_ok:
            raise ValueError("the fused backend needs CUDA, label_size<=256, n_components<=236 and hidden sizes<=256")
        self.backend = backend
        self.engine = None
        self._pending_opt_state = None

        if backend == "fused":
            self.optimizer = FusedAdamHandle(self)
        else:
            self.optimizer = self._make_torch_optimizer()
        if self.reduce_on_plateau:
            # fused backend: the same rule (torch defaults: factor 0.1, relative threshold 1e-4) on the engine's lr
            self.scheduler = (PlateauLR(self, patience=10) if backend == "fused"
                              else ReduceLROnPlateau(self.optimizer, patience=10))

    # ------------------------------------------------------------------------------------------
    def _make_torch_optimizer(self):
        ps = self.model.parameters()
        if self.solver == "adam":
            return optim.Adam(ps, lr=self.lr, betas=(self.momentum, 0.99))
        if self.solver == "sgd":
            return optim.SGD(ps, lr=self.lr, momentum=self.momentum)
        if self.solver == "adagrad":
            return optim.Adagrad(ps, lr=self.lr)
        if self.solver == "adadelta":
            return optim.Adadelta(ps, lr=self.lr)
        return optim.RMSprop(ps, lr=self.lr, momentum=self.momentum)

    def _ensure_engine(self, loader: CsrMinibatchLoader) -> None:
        if self.backend != "fused":
            return
        need_new = (self.engine is None or loader.batch_size > self.engine.max_batch
                    or loader.max_nnz > self.engine.max_nnz)
        if not need_new:
            return
        from ..engine.fused import FusedEngine

        old = self.engine
        self.engine = FusedEngine(self.model, max_batch=loader.batch_size, max_nnz=max(loader.max_nnz, 1),
                                  lr=self.lr, betas=(self.momentum, 0.99), kl_weight=float(self.weights["beta"]),
                                  seed=self.seed, device=self.device, solver=self.solver)
        if old is not None:  # keep optimiser state across a workspace re-allocation
            self.engine.flat.M.copy_(old.flat.M)
            self.engine.flat.V2.copy_(old.flat.V2)
            self.engine.set_step_count(old.step_count)
            self.engine.ctr[0] = old.ctr[0]
            self.engine.rng_calls = old.rng_calls
        self._apply_pending_opt_state()

    def _apply_pending_opt_state(self) -> None:
        if self.engine is None or not self._pending_opt_state:
            return
        eng = self.engine
        names = [n for n, _ in self.model.named_parameters()]
        step = 0
        for idx, st in self._pending_opt_state.items():
            name = names[int(idx)]
            eng.flat.view(eng.flat.M, name).copy_(torch.as_tensor(st["exp_avg"]).to(eng.device))
            eng.flat.view(eng.flat.V2, name).copy_(torch.as_tensor(st["exp_avg_sq"]).to(eng.device))
            step = max(step, int(float(torch.as_tensor(st["step"]).reshape(-1)[0])))
        eng.set_step_count(step)
        self._pending_opt_state = None

    def make_loader(self, dataset, shuffle: bool, batch_size: Optional[int] = None) -> CsrMinibatchLoader:
        return CsrMinibatchLoader(dataset, batch_size or self.batch_size, shuffle=shuffle, device=self.device,
                                  seed=self.seed + (0 if shuffle else 1))

    def parameters_changed(self) -> None:
        """Call after writing parameters from outside (``load_state_dict``, federated averaging)."""
        if self.engine is not None:
            self.engine.sync_shadows()

    # ------------------------------------------------------------------------------------------
    # loss
    # ------------------------------------------------------------------------------------------
    def _loss_terms(self, inputs, word_dists, prior_mean, prior_variance, posterior_mean, posterior_variance,
                    posterior_log_variance):
        return elbo_terms(inputs, word_dists, prior_mean, prior_variance, posterior_mean, posterior_variance,
                          posterior_log_variance)

    def _eager_forward_loss(self, mb: Minibatch):
        x = mb.bow.to_dense()
        if self._family == "ctm" or self.infnet != "bow":
            out = self.model(x, mb.x_ctx, mb.labels)
            pm, pv, mu, var, logvar, wd, est = out
        else:
            pm, pv, mu, var, logvar, wd = self.model(x)
            est = None
        kl, rl = self._loss_terms(x, wd, pm, pv, mu, var, logvar)
        loss = (self.weights["beta"] * kl + rl).sum()
        if est is not None and mb.labels is not None:
            target = torch.argmax(mb.labels, 1)
            loss = loss + torch.nn.CrossEntropyLoss()(est, target)
        return loss

    def train_minibatch(self, mb: Minibatch) -> torch.Tensor:
        """One optimisation step on a staged minibatch; returns the summed loss (0-d/1-d tensor,
        still on the device)."""
        if self.backend == "fused":
            return self.engine.train_step(mb.bow, mb.x_ctx, labels=mb.labels).clone()
        self.model.zero_grad()
        loss = self._eager_forward_loss(mb)
        loss.backward()
        self.optimizer.step()
        return loss.detach()

    def _train_epoch(self, loader):
        self.model.train()
        self._ensure_engine(loader)
        total = None
        samples = 0
        for mb in loader:
            loss = self.train_minibatch(mb)
            total = loss.double().reshape(()) if total is None else total + loss.double().reshape(())
            samples += mb.size
        train_loss = float(total.item()) / max(samples, 1) if total is not None else 0.0
        return samples, train_loss

    def _validate_epoch(self, loader):
        self.model.eval()
        self._ensure_engine(loader)
        total = 0.0
        samples = 0
        with torch.no_grad():
            for mb in loader:
                if self.backend == "fused":
                    loss = self.engine.forward_backward(mb.bow, mb.x_ctx, training=False, need_grad=False, labels=mb.labels)
                    total += float(loss.item())
                else:
                    total += float(self._eager_forward_loss(mb).item())
                samples += mb.size
        return samples, total / max(samples, 1)

    # ------------------------------------------------------------------------------------------
    # fit
    # ------------------------------------------------------------------------------------------
    def fit(self, train_dataset, validation_dataset=None, save_dir=None, patience=5, delta=0, n_samples=20):
        if self.verbose:
            print("Settings: \n               N Components: {}\n               Topic Prior Mean: {}\n"
                  "               Topic Prior Variance: {}\n               Model Type: {}\n"
                  "               Hidden Sizes: {}\n               Activation: {}\n               Dropout: {}\n"
                  "               Learn Priors: {}\n               Learning Rate: {}\n               Momentum: {}\n"
                  "               Reduce On Plateau: {}\n               Save Dir: {}".format(
                      self.n_components, self.topic_prior_mean, self.topic_prior_variance, self.model_type,
                      self.hidden_sizes, self.activation, self.dropout, self.learn_priors, self.lr,
                      self.momentum, self.reduce_on_plateau, save_dir))
        self.model_dir = save_dir
        self.train_data = train_dataset
        self.validation_data = validation_dataset
        if self.validation_data is not None:
            self.early_stopping = EarlyStopping(patience=patience, verbose=self.verbose, path=save_dir, delta=delta)
        train_loader = self.make_loader(train_dataset, shuffle=True)
        val_loader = self.make_loader(validation_dataset, shuffle=True) if validation_dataset is not None else None
        samples_processed = 0
        for epoch in range(self.num_epochs):
            self.nn_epoch = epoch
            s = datetime.datetime.now()
            sp, train_loss = self._train_epoch(train_loader)
            samples_processed += sp
            e = datetime.datetime.now()
            self.best_components = self.model.beta
            if val_loader is not None:
                _, val_loss = self._validate_epoch(val_loader)
                if self.verbose:
                    self.logger.info("Epoch: [{}/{}]\tSamples: [{}/{}]\tTrain Loss: {}\tValid Loss: {}\tTime: {}".format(
                        epoch + 1, self.num_epochs, samples_processed, len(train_dataset) * self.num_epochs,
                        train_loss, val_loss, e - s))
                if np.isnan(val_loss) or np.isnan(train_loss):
                    break
                if self.reduce_on_plateau:
                    self.scheduler.step(val_loss)
                self.early_stopping(val_loss, self)
                if self.early_stopping.early_stop:
                    self.logger.info("Early stopping")
                    break
            else:
                if save_dir is not None:
                    self.save(save_dir)
                if self.verbose:
                    self.logger.info("Epoch: [{}/{}]\tSamples: [{}/{}]\tTrain Loss: {}\tTime: {}".format(
                        epoch + 1, self.num_epochs, samples_processed, len(train_dataset) * self.num_epochs,
                        train_loss, e - s))
            if train_loss < self.best_loss_train:
                self.best_loss_train = train_loss
        self.training_doc_topic_distributions = self.get_doc_topic_distribution(train_dataset, n_samples)

    # ------------------------------------------------------------------------------------------
    # inference / results
    # ------------------------------------------------------------------------------------------
    def get_doc_topic_distribution(self, dataset, n_samples=20, thr: float = 0.0):
        """Mean of ``n_samples`` stochastic theta draws per document, ``[n_docs, K]`` numpy array.  ``thr > 0``:
        entries below it are zeroed and the rows L1-renormalised (federated_model.py:170-173), on the device for
        the fused backend."""
        self.model.eval()
        loader = self.make_loader(dataset, shuffle=False)
        self._ensure_engine(loader)
        out = np.zeros((len(dataset), self.n_components), dtype=np.float64)
        with torch.no_grad():
            for mb in loader:
                if self.backend == "fused":
                    th = self.engine.sample_theta(mb.bow, mb.x_ctx, n_samples=n_samples, thr=thr, labels=mb.labels)
                else:
                    x = mb.bow.to_dense()
                    th = torch.zeros(mb.size, self.n_components, device=self.device)
                    for _ in range(n_samples):
                        th += self.model.get_theta(x, mb.x_ctx, mb.labels)
                    th /= n_samples
                out[mb.index] = th.double().cpu().numpy()
        if thr > 0 and self.backend != "fused":
            from ..utils.metrics import thresholded_thetas

            out = thresholded_thetas(out, thr)
        return out

    def get_thetas(self, dataset, n_samples=20):
        return self.get_doc_topic_distribution(dataset, n_samples=n_samples)

    def get_predicted_topics(self, dataset, n_samples):
        thetas = self.get_doc_topic_distribution(dataset, n_samples)
        return [int(np.argmax(thetas[i] / np.sum(thetas[i]))) for i in range(len(dataset))]

    def get_most_likely_topic(self, doc_topic_distribution):
        return np.argmax(doc_topic_distribution, axis=0)

    def _current_topic_word_matrix(self) -> torch.Tensor:
        m = self.model
        if m.model_type.lower() == "prodlda":
            return m.beta.detach()
        with torch.no_grad():
            bn = torch.nn.functional.batch_norm(m.beta, m.beta_batchnorm.running_mean, m.beta_batchnorm.running_var,
                                                training=False, eps=m.beta_batchnorm.eps)
            return torch.softmax(bn, dim=1)

    def get_topic_word_matrix(self):
        """ProdLDA: beta (un-normalised); LDA: the simplex-constrained matrix."""
        return self._current_topic_word_matrix().cpu().numpy()

    def get_topic_word_distribution(self):
        return softmax(self.get_topic_word_matrix(), axis=1)

    def get_word_distribution_by_topic_id(self, topic):
        if topic >= self.n_components:
            raise Exception("Topic id must be lower than the number of topics")
        wd = self.get_topic_word_distribution()
        t = [(word, wd[topic][idx]) for idx, word in self._idx2token_items()]
        return sorted(t, key=lambda x: -x[1])

    def _idx2token_items(self):
        i2t = self.train_data.idx2token
        if isinstance(i2t, dict):
            return sorted(i2t.items())
        return list(enumerate(i2t))

    def _idx2token(self, idx: int) -> str:
        return self.train_data.idx2token[idx]

    def get_topics(self, k=10) -> List[List[str]]:
        """Top-``k`` words of every topic (``best_components``), via the row top-k kernel on GPU."""
        assert k <= self.input_size, "k must be <= input size."
        comp = self.best_components if self.best_components is not None else self.model.beta
        comp = comp.detach()
        if comp.is_cuda and k <= 32:
            from ..ops import kernels as K_

            _, idxs = K_.row_topk(comp, k)
        else:
            _, idxs = torch.topk(comp, k, dim=1)
        idxs = idxs.cpu().numpy()
        return [[self._idx2token(int(j)) for j in idxs[i]] for i in range(self.n_components)]

    def get_topic_lists(self, k=10):
        return self.get_topics(k)

    def get_top_documents_per_topic_id(self, unpreprocessed_corpus, document_topic_distributions, topic_id, k=5):
        probability_list = document_topic_distributions.T[topic_id]
        ind = probability_list.argsort()[-k:][::-1]
        return [(unpreprocessed_corpus[i], document_topic_distributions[i][topic_id]) for i in ind]

    def get_ldavis_data_format(self, vocab, dataset, n_samples):
        term_frequency = np.ravel(dataset.X.sum(axis=0))
        doc_lengths = np.ravel(dataset.X.sum(axis=1))
        return {"topic_term_dists": self.get_topic_word_distribution(),
                "doc_topic_dists": self.get_doc_topic_distribution(dataset, n_samples),
                "doc_lengths": doc_lengths, "vocab": vocab, "term_frequency": term_frequency}

    # ------------------------------------------------------------------------------------------
    # checkpoints
    # ------------------------------------------------------------------------------------------
    _file_prefix = "AVITM"

    def _format_file(self):
        # the reference's directory name shifts labels vs values (avitm.py:591-596); kept verbatim so
        # directories written by either implementation are found by the other
        return "{}_nc_{}_tpm_{}_tpv_{}_hs_{}_ac_{}_do_{}_lr_{}_mo_{}_rp_{}".format(
            self._file_prefix, self.n_components, 0.0, 1 - (1. / self.n_components), self.model_type,
            self.hidden_sizes, self.activation, self.dropout, self.lr, self.momentum, self.reduce_on_plateau)

    def _checkpoint_attrs(self) -> dict:
        skip = {"model", "engine", "optimizer", "scheduler", "logger", "early_stopping", "train_data",
                "validation_data", "best_components", "training_doc_topic_distributions", "_pending_opt_state"}
        return {k: v for k, v in self.__dict__.items() if k not in skip}

    def save(self, models_dir=None):
        """``<models_dir>/<_format_file()>/epoch_<n>.pth`` with keys ``state_dict`` and ``dcue_dict``
        (plus ``optimizer`` so training can resume).  Ref: avitm.py:598-617."""
        if self.model is None or models_dir is None:
            return
        d = os.path.join(models_dir, self._format_file())
        os.makedirs(d, exist_ok=True)
        fileloc = os.path.join(d, "epoch_{}.pth".format(self.nn_epoch))
        sd = {k: v.detach().cpu().contiguous() for k, v in self.model.state_dict().items()}
        with open(fileloc, "wb") as f:
            ck = {"state_dict": sd, "dcue_dict": self._checkpoint_attrs(), "optimizer": self.optimizer.state_dict()}
            td = getattr(self, "train_data", None)
            if td is not None and getattr(td, "idx2token", None) is not None:
                # extra key (ignored by the reference's loader): lets hold-out inference rebuild the vocabulary
                i2t = td.idx2token
                ck["vocab"] = [i2t[i] for i in range(len(i2t))]
            torch.save(ck, f)

    def load(self, model_dir, epoch):
        """Restores a checkpoint written by :meth:`save`.  Ref: avitm.py:619-640."""
        model_file = os.path.join(model_dir, "epoch_" + str(epoch) + ".pth")
        with open(model_file, "rb") as f:
            checkpoint = torch.load(f, map_location="cpu", weights_only=False)
        for k, v in checkpoint["dcue_dict"].items():
            if k not in ("device", "USE_CUDA", "backend"):
                setattr(self, k, v)
        self.model.load_state_dict(checkpoint["state_dict"])
        self.parameters_changed()
        if "optimizer" in checkpoint and checkpoint["optimizer"].get("state"):
            try:
                self.optimizer.load_state_dict(checkpoint["optimizer"])
            except Exception as ex:  # a checkpoint written by another solver
                self.logger.warning("optimizer state not restored: %s", ex)
