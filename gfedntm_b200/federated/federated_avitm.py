"""Federated AVITM (ProdLDA / NeuralLDA).  Ref: src/models/federated/federated_avitm.py:14-193."""
from __future__ import annotations

from ..models.avitm import AVITM
from .federated_model import FederatedModel


class FederatedAVITM(AVITM, FederatedModel):
    def __init__(self, tm_params: dict, grads_to_share, logger=None, share: str = "param", backend: str = "auto",
                 device=None) -> None:
        FederatedModel.__init__(self, tm_params, grads_to_share, logger, share=share)
        AVITM.__init__(self, logger=self.logger, input_size=tm_params["input_size"],
                       n_components=tm_params["n_components"], model_type=tm_params["model_type"],
                       hidden_sizes=tm_params["hidden_sizes"], activation=tm_params["activation"],
                       dropout=tm_params["dropout"], learn_priors=tm_params["learn_priors"],
                       batch_size=tm_params["batch_size"], lr=tm_params["lr"], momentum=tm_params["momentum"],
                       solver=tm_params["solver"], num_epochs=tm_params["num_epochs"],
                       reduce_on_plateau=tm_params["reduce_on_plateau"],
                       topic_prior_mean=tm_params["topic_prior_mean"],
                       topic_prior_variance=tm_params["topic_prior_variance"],
                       num_samples=tm_params["num_samples"],
                       num_data_loader_workers=tm_params["num_data_loader_workers"],
                       verbose=tm_params["verbose"], backend=backend, device=device,
                       seed=tm_params.get("seed"))
