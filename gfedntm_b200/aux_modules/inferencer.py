"""Hold-out inference from saved models (the role of the submodule's inferencers,
src/aux_modules/topicmodeler/src/topicmodeling/inferencer.py:17-400): rebuild a trained ProdLDA / CTM from a
checkpoint directory written by ``model.save()``, vectorise new documents against the TRAINING vocabulary and
write their topic distributions as ``doc-topics.npy`` (+ ``doc-topics.csv``).

The training vocabulary travels with the checkpoint (``model.save()`` adds a ``vocab`` list next to ``state_dict`` /
``dcue_dict`` when the training dataset has one); it can also be passed explicitly.
"""
from __future__ import annotations

import glob
import os
import re
from typing import Optional, Sequence

import numpy as np
import scipy.sparse as sp
import torch
from sklearn.feature_extraction.text import CountVectorizer

from ..models import AVITM, CTM, BOWDataset, CTMDataset


def latest_epoch(model_dir: str) -> int:
    eps = [int(m.group(1)) for f in glob.glob(os.path.join(model_dir, "epoch_*.pth"))
           if (m := re.search(r"epoch_(\d+)\.pth$", f))]
    if not eps:
        raise FileNotFoundError(f"no epoch_<n>.pth checkpoint in {model_dir}")
    return max(eps)


class Inferencer:
    """``Inferencer(model_dir).predict(texts)`` -> thetas ``[n_docs, K]``."""

    def __init__(self, model_dir: str, epoch: Optional[int] = None, vocab: Optional[Sequence[str]] = None,
                 device: Optional[str] = None, thetas_thr: float = 3e-3) -> None:
        self.model_dir = model_dir
        self.epoch = latest_epoch(model_dir) if epoch is None else int(epoch)
        with open(os.path.join(model_dir, f"epoch_{self.epoch}.pth"), "rb") as f:
            ck = torch.load(f, map_location="cpu", weights_only=False)
        a = ck["dcue_dict"]
        self.is_ctm = "contextual_size" in a and a.get("contextual_size")
        common = dict(input_size=a["input_size"], n_components=a["n_components"], model_type=a["model_type"],
                      hidden_sizes=tuple(a["hidden_sizes"]), activation=a["activation"], dropout=a["dropout"],
                      learn_priors=a.get("learn_priors", True), batch_size=a.get("batch_size", 64),
                      lr=a.get("lr", 2e-3), momentum=a.get("momentum", 0.99), verbose=False,
                      device=device or ("cuda" if torch.cuda.is_available() else "cpu"))
        if self.is_ctm:
            self.model = CTM(contextual_size=a["contextual_size"], inference_type=a.get("inference_type", "combined"),
                             **common)
        else:
            self.model = AVITM(**common)
        self.model.load(model_dir, self.epoch)
        vocab = ck.get("vocab") if vocab is None else vocab
        if vocab is None:
            raise ValueError("the checkpoint carries no vocabulary: pass `vocab`")
        self.vocab = [str(t) for t in vocab]
        if len(self.vocab) != a["input_size"]:
            raise ValueError("vocabulary size does not match the model's input size")
        self.thr = thetas_thr

    def vectorise(self, texts: Sequence[str]) -> sp.csr_matrix:
        cv = CountVectorizer(vocabulary={t: i for i, t in enumerate(self.vocab)}, lowercase=True)
        return cv.transform(texts).astype(np.float32)

    def predict(self, texts: Optional[Sequence[str]] = None, bow: Optional[sp.spmatrix] = None,
                embeddings: Optional[np.ndarray] = None, n_samples: int = 20, out_dir: Optional[str] = None
                ) -> np.ndarray:
        if bow is None:
            if texts is None:
                raise ValueError("give `texts` or a document-term matrix `bow`")
            bow = self.vectorise(texts)
        if self.is_ctm:
            if embeddings is None:
                raise ValueError("a contextualized model needs `embeddings` for the hold-out documents")
            ds = CTMDataset(np.asarray(embeddings, dtype=np.float32), bow, self.vocab)
        else:
            ds = BOWDataset(bow, self.vocab)
        thetas = np.asarray(self.model.get_doc_topic_distribution(ds, n_samples=n_samples), dtype=np.float64)
        thetas[thetas < self.thr] = 0.0
        s = thetas.sum(1, keepdims=True)
        s[s == 0] = 1.0
        thetas = thetas / s
        if out_dir is not None:
            os.makedirs(out_dir, exist_ok=True)
            np.save(os.path.join(out_dir, "doc-topics.npy"), thetas)
            np.savetxt(os.path.join(out_dir, "doc-topics.csv"), thetas, delimiter=",", fmt="%.6f")
        return thetas
