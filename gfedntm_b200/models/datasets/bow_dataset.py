"""Bag-of-words dataset.  Ref: src/models/base/pytorchavitm/datasets/bow_dataset.py:6-34.

The reference keeps a dense ``[n_docs, V]`` float matrix on the host.  Here the corpus is stored
once as CSR (``scipy.sparse.csr_matrix``, int32 indices / float32 counts); ``__getitem__`` still
returns ``{'X': FloatTensor[V]}`` so user code written against the reference keeps working,
while the training loops read CSR minibatches directly (``csr_rows``).
"""
from __future__ import annotations

from typing import Dict, Sequence

import numpy as np
import scipy.sparse as sp
import torch
from torch.utils.data import Dataset


def to_csr32(X) -> sp.csr_matrix:
    """Dense array / torch tensor / any scipy sparse matrix -> canonical CSR (int32 / float32)."""
    if isinstance(X, torch.Tensor):
        X = X.detach().cpu().numpy()
    if sp.issparse(X):
        m = X.tocsr()
    else:
        m = sp.csr_matrix(np.asarray(X))
    m = m.astype(np.float32)
    m.sum_duplicates()
    m.sort_indices()
    m.indptr = m.indptr.astype(np.int32)
    m.indices = m.indices.astype(np.int32)
    return m


class BOWDataset(Dataset):
    """Documents as bag-of-words count vectors plus the index -> token map."""

    def __init__(self, X, idx2token) -> None:
        self.X = to_csr32(X)
        self.idx2token = idx2token
        if isinstance(idx2token, dict):
            n_tok = len(idx2token)
        else:
            n_tok = len(idx2token) if idx2token is not None else self.X.shape[1]
        if idx2token is not None and n_tok != self.X.shape[1]:
            raise ValueError(f"idx2token has {n_tok} entries but X has {self.X.shape[1]} columns")

    def __len__(self) -> int:
        return self.X.shape[0]

    def __getitem__(self, i: int) -> Dict[str, torch.Tensor]:
        row = np.asarray(self.X[i].todense(), dtype=np.float32).reshape(-1)
        return {"X": torch.from_numpy(row)}

    @property
    def vocab_size(self) -> int:
        return self.X.shape[1]

    def csr_rows(self, idx: Sequence[int]) -> sp.csr_matrix:
        return self.X[np.asarray(idx)]
