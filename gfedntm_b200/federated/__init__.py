from .federated_avitm import FederatedAVITM
from .federated_ctm import FederatedCTM
from .federated_model import FederatedModel

__all__ = ["FederatedModel", "FederatedAVITM", "FederatedCTM"]
