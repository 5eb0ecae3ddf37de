from .bow_dataset import BOWDataset
from .ctm_dataset import CTMDataset

__all__ = ["BOWDataset", "CTMDataset"]
