"""Contextualized topic models: CTM, ZeroShotTM, CombinedTM.
Ref: src/models/base/contextualized_topic_models/ctm_network/ctm.py:20-807."""
from __future__ import annotations

import multiprocessing as mp

from .base import NeuralTopicModel


class CTM(NeuralTopicModel):
    _family = "ctm"
    _file_prefix = "contextualized_topic_model"

    def __init__(self, logger=None, input_size=None, contextual_size=None, inference_type="combined",
                 n_components=10, model_type="prodLDA", hidden_sizes=(100, 100), activation="softplus",
                 dropout=0.2, learn_priors=True, batch_size=64, lr=2e-3, momentum=0.99, solver="adam",
                 num_epochs=100, num_samples=10, reduce_on_plateau=False, topic_prior_mean=0.0,
                 topic_prior_variance=None, num_data_loader_workers=mp.cpu_count(), label_size=0,
                 loss_weights=None, verbose=True, backend="auto", device=None, seed=None):
        if inference_type not in ("combined", "zeroshot"):
            raise Exception("Missing infnet parameter, options are zeroshot and combined")
        assert isinstance(contextual_size, int) and contextual_size > 0, "contextual_size must be int > 0"
        self.inference_type = inference_type
        self._setup(logger=logger, input_size=input_size, n_components=n_components, model_type=model_type,
                    hidden_sizes=hidden_sizes, activation=activation, dropout=dropout,
                    learn_priors=learn_priors, batch_size=batch_size, lr=lr, momentum=momentum,
                    solver=solver, num_epochs=num_epochs, reduce_on_plateau=reduce_on_plateau,
                    topic_prior_mean=topic_prior_mean, topic_prior_variance=topic_prior_variance,
                    num_samples=num_samples, num_data_loader_workers=num_data_loader_workers,
                    verbose=verbose, infnet=inference_type, contextual_size=contextual_size,
                    label_size=label_size, loss_weights=loss_weights, backend=backend, device=device,
                    seed=seed)

    def _loss(self, inputs, word_dists, prior_mean, prior_variance, posterior_mean, posterior_variance,
              posterior_log_variance):
        """Per-document (KL, RL).  Ref: ctm.py:182-238."""
        return self._loss_terms(inputs, word_dists, prior_mean, prior_variance, posterior_mean,
                                posterior_variance, posterior_log_variance)


class ZeroShotTM(CTM):
    """CTM whose encoder sees only the contextual embedding.  Ref: ctm.py:785-798."""

    def __init__(self, **kwargs):
        kwargs["inference_type"] = "zeroshot"
        super().__init__(**kwargs)


class CombinedTM(CTM):
    """CTM whose encoder sees BoW + adapted contextual embedding.  Ref: ctm.py:801-807."""

    def __init__(self, **kwargs):
        kwargs["inference_type"] = "combined"
        super().__init__(**kwargs)
