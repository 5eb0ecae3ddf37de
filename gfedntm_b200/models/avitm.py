"""AVITM trainer (ProdLDA / NeuralLDA).
Ref: src/models/base/pytorchavitm/avitm_network/avitm.py:20-640 — same constructor signature,
defaults, public methods and checkpoint naming; execution goes through the fused sm_100a engine on
CUDA devices (see :mod:`gfedntm_b200.models.base`)."""
from __future__ import annotations

from .base import NeuralTopicModel


class AVITM(NeuralTopicModel):
    _family = "avitm"
    _file_prefix = "AVITM"

    def __init__(self, logger=None, input_size=None, n_components=10, model_type="prodLDA",
                 hidden_sizes=(100, 100), activation="softplus", dropout=0.2, learn_priors=True,
                 batch_size=64, lr=2e-3, momentum=0.99, solver="adam", num_epochs=100,
                 reduce_on_plateau=False, topic_prior_mean=0.0, topic_prior_variance=None,
                 num_samples=10, num_data_loader_workers=0, verbose=True, backend="auto", device=None,
                 seed=None):
        self._setup(logger=logger, input_size=input_size, n_components=n_components, model_type=model_type,
                    hidden_sizes=hidden_sizes, activation=activation, dropout=dropout,
                    learn_priors=learn_priors, batch_size=batch_size, lr=lr, momentum=momentum,
                    solver=solver, num_epochs=num_epochs, reduce_on_plateau=reduce_on_plateau,
                    topic_prior_mean=topic_prior_mean, topic_prior_variance=topic_prior_variance,
                    num_samples=num_samples, num_data_loader_workers=num_data_loader_workers,
                    verbose=verbose, infnet="bow", backend=backend, device=device, seed=seed)

    def _loss(self, inputs, word_dists, prior_mean, prior_variance, posterior_mean, posterior_variance,
              posterior_log_variance):
        """Summed ELBO loss of a batch.  Ref: avitm.py:168-229."""
        kl, rl = self._loss_terms(inputs, word_dists, prior_mean, prior_variance, posterior_mean,
                                  posterior_variance, posterior_log_variance)
        return (kl + rl).sum()
