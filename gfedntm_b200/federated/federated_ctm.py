"""Federated CTM.  Ref: src/models/federated/federated_ctm.py:14-190.

The reference never forwards the inference type, so its federated CTM is always CombinedTM; here the
``ctm_model_type`` configuration key (``CombinedTM`` / ``ZeroShotTM``) is honoured."""
from __future__ import annotations

from ..models.ctm import CTM
from .federated_model import FederatedModel


class FederatedCTM(CTM, FederatedModel):
    def __init__(self, tm_params: dict, grads_to_share, logger=None, share: str = "param", backend: str = "auto",
                 device=None) -> None:
        FederatedModel.__init__(self, tm_params, grads_to_share, logger, share=share)
        kind = str(tm_params.get("ctm_model_type", "CombinedTM")).lower()
        inference_type = "zeroshot" if kind.startswith("zero") else "combined"
        CTM.__init__(self, logger=self.logger, input_size=tm_params["input_size"],
                     contextual_size=tm_params["contextual_size"], inference_type=inference_type,
                     n_components=tm_params["n_components"], model_type=tm_params["model_type"],
                     hidden_sizes=tm_params["hidden_sizes"], activation=tm_params["activation"],
                     dropout=tm_params["dropout"], learn_priors=tm_params["learn_priors"],
                     batch_size=tm_params["batch_size"], lr=tm_params["lr"], momentum=tm_params["momentum"],
                     solver=tm_params["solver"], num_epochs=tm_params["num_epochs"],
                     num_samples=tm_params["num_samples"], reduce_on_plateau=tm_params["reduce_on_plateau"],
                     topic_prior_mean=tm_params["topic_prior_mean"],
                     topic_prior_variance=tm_params["topic_prior_variance"],
                     num_data_loader_workers=tm_params["num_data_loader_workers"],
                     label_size=int(tm_params.get("label_size", 0)), verbose=tm_params["verbose"],
                     backend=backend, device=device, seed=tm_params.get("seed"))

    def _epoch_end(self, save_dir) -> None:
        """Full result snapshot at every epoch end, ``<save_dir>_epoch_<n>`` (federated_ctm.py:150-159)."""
        if not save_dir:
            return
        try:
            self.logger.info(f"-- -- Saving model at epoch {self.current_epoch + 1}")
            self.get_results_model(save_dir + f"_epoch_{self.current_epoch + 1}")
        except Exception as ex:  # keep training if the snapshot fails, like the reference
            self.logger.info(f"-- -- Error while saving model at epoch {self.current_epoch + 1}: {ex}")
