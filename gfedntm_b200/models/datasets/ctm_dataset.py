"""Dataset for contextualized topic models.
Ref: src/models/base/contextualized_topic_models/datasets/dataset.py:6-48."""
from __future__ import annotations

from typing import Dict, Optional, Sequence

import numpy as np
import scipy.sparse as sp
import torch
from torch.utils.data import Dataset

from .bow_dataset import to_csr32


class CTMDataset(Dataset):
    """Bag of words + a fixed-size contextual embedding per document (+ optional one-hot labels)."""

    def __init__(self, X_contextual, X_bow, idx2token, qt=None, labels=None) -> None:
        # `qt`: opaque preparation object the reference stores on the dataset (datasets/dataset.py:9,24);
        # fourth positional argument there, optional here
        self.qt = qt
        self.X_bow = to_csr32(X_bow)
        self.X_contextual = np.ascontiguousarray(np.asarray(X_contextual, dtype=np.float32))
        if self.X_bow.shape[0] != len(self.X_contextual):
            raise Exception("Wait! BoW and Contextual Embeddings have different sizes! "
                            "You might want to check if the BoW preparation method has removed some documents.")
        if labels is not None:
            labels = labels if sp.issparse(labels) else np.asarray(labels, dtype=np.float32)
            if labels.shape[0] != self.X_bow.shape[0]:
                raise Exception(f"There is something wrong in the length of the labels (size: {labels.shape[0]}) "
                                f"and the bow (len: {self.X_bow.shape[0]}). These two numbers should match.")
        self.labels = labels
        self.idx2token = idx2token

    def __len__(self) -> int:
        return self.X_bow.shape[0]

    def __getitem__(self, i: int) -> Dict[str, torch.Tensor]:
        row = np.asarray(self.X_bow[i].todense(), dtype=np.float32).reshape(-1)
        item = {"X_bow": torch.from_numpy(row), "X_contextual": torch.from_numpy(self.X_contextual[i])}
        if self.labels is not None:
            lab = self.labels[i]
            lab = np.asarray(lab.todense() if sp.issparse(lab) else lab, dtype=np.float32).reshape(-1)
            item["labels"] = torch.from_numpy(lab)
        return item

    @property
    def X(self) -> sp.csr_matrix:  # same accessor as BOWDataset for the shared loaders
        return self.X_bow

    @property
    def vocab_size(self) -> int:
        return self.X_bow.shape[1]

    def label_rows(self, idx: Sequence[int]) -> Optional[np.ndarray]:
        if self.labels is None:
            return None
        lab = self.labels[np.asarray(idx)]
        return np.asarray(lab.todense() if sp.issparse(lab) else lab, dtype=np.float32)
