from .avitm import AVITM
from .ctm import CTM, CombinedTM, ZeroShotTM
from .datasets import BOWDataset, CTMDataset
from .early_stopping import EarlyStopping
from .networks import DecoderNetwork, InferenceNetwork

__all__ = ["AVITM", "CTM", "CombinedTM", "ZeroShotTM", "BOWDataset", "CTMDataset", "EarlyStopping",
           "DecoderNetwork", "InferenceNetwork"]
