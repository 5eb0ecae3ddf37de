"""Persisted topic model with curation operations (the role of the submodule's ``TMmodel``,
src/aux_modules/topicmodeler/src/topicmodeling/manageModels.py:321-1424).

A model folder holds the matrices that characterise a trained model and the quantities derived from them,
under the submodule's file names so folders are interchangeable:

    alphas.npy  betas.npy  thetas.npz  vocab.txt            current state
    alphas_orig.npy  betas_orig.npy  thetas_orig.npz         state at creation (``resetTM``)
    betas_ds.npy  topic_entropy.npy  topic_coherence.npy  ndocs_active.npy
    tpc_descriptions.txt  tpc_labels.txt  edits.txt

Design difference: every derived quantity is a pure function of ``(betas, thetas)``, recomputed by one
``_derive`` call after each edit, instead of being patched per operation; the coherence is NPMI on a
document-term matrix supplied by the caller (``gfedntm_b200.utils.metrics``), not a gensim ``CoherenceModel``.
"""
from __future__ import annotations

import logging
from pathlib import Path
from typing import List, Optional, Sequence

import numpy as np
import scipy.sparse as sp

from ..utils import metrics


class TMmodel:
    def __init__(self, TMfolder, logger: Optional[logging.Logger] = None, thetas_thr: float = 3e-3) -> None:
        self._TMfolder = Path(TMfolder)
        self._logger = logger or logging.getLogger("TMmodel")
        self._thr = float(thetas_thr)
        self._alphas = self._betas = self._thetas = None
        self._vocab: List[str] = []
        self._edits: List[str] = []
        self._tpc_labels: List[str] = []
        self._ref_bow = None
        if (self._TMfolder / "betas.npy").exists():
            self._load()

    # ------------------------------------------------------------------ creation / persistence
    def create(self, betas, thetas, alphas=None, vocab: Sequence[str] = (), labels: Optional[Sequence[str]] = None,
               reference_bow=None) -> None:
        """``betas`` [K,V], ``thetas`` [D,K] (dense or sparse; entries below ``thetas_thr`` are dropped and the
        rows renormalised, as the trainers do), ``alphas`` [K] topic sizes (default: mean of thetas)."""
        self._TMfolder.mkdir(parents=True, exist_ok=True)
        betas = np.asarray(betas, dtype=np.float64)
        th = np.asarray(thetas.todense() if sp.issparse(thetas) else thetas, dtype=np.float64)
        th = metrics.thresholded_thetas(th, self._thr) if hasattr(metrics, "thresholded_thetas") else th
        th = sp.csr_matrix(th)
        alphas = np.asarray(th.mean(0)).ravel() if alphas is None else np.asarray(alphas, dtype=np.float64)
        self._betas, self._thetas, self._alphas = betas, th, alphas
        self._vocab = list(vocab)
        self._edits = []
        self._ref_bow = reference_bow
        np.save(self._TMfolder / "alphas_orig.npy", alphas)
        np.save(self._TMfolder / "betas_orig.npy", betas)
        sp.save_npz(self._TMfolder / "thetas_orig.npz", th)
        (self._TMfolder / "vocab.txt").write_text("\n".join(self._vocab), encoding="utf8")
        self._tpc_labels = list(labels) if labels is not None else [f"topic {k}" for k in range(betas.shape[0])]
        self._sort_topics()
        self._derive()
        self._save_all()

    def _save_all(self) -> None:
        f = self._TMfolder
        np.save(f / "alphas.npy", self._alphas)
        np.save(f / "betas.npy", self._betas)
        sp.save_npz(f / "thetas.npz", self._thetas)
        np.save(f / "betas_ds.npy", self._betas_ds)
        np.save(f / "topic_entropy.npy", self._topic_entropy)
        np.save(f / "topic_coherence.npy", self._topic_coherence)
        np.save(f / "ndocs_active.npy", self._ndocs_active)
        (f / "tpc_descriptions.txt").write_text("\n".join(self._tpc_descriptions), encoding="utf8")
        (f / "tpc_labels.txt").write_text("\n".join(self._tpc_labels), encoding="utf8")
        (f / "edits.txt").write_text("\n".join(self._edits), encoding="utf8")

    def _load(self) -> None:
        f = self._TMfolder
        self._alphas = np.load(f / "alphas.npy")
        self._betas = np.load(f / "betas.npy")
        self._thetas = sp.load_npz(f / "thetas.npz").tocsr()
        self._vocab = (f / "vocab.txt").read_text(encoding="utf8").split("\n")
        self._betas_ds = np.load(f / "betas_ds.npy")
        self._topic_entropy = np.load(f / "topic_entropy.npy")
        self._topic_coherence = np.load(f / "topic_coherence.npy")
        self._ndocs_active = np.load(f / "ndocs_active.npy")
        rd = lambda n: [l for l in (f / n).read_text(encoding="utf8").split("\n") if l] if (f / n).exists() else []  # noqa: E731,E741
        self._tpc_descriptions, self._tpc_labels, self._edits = rd("tpc_descriptions.txt"), rd("tpc_labels.txt"), rd("edits.txt")

    # ------------------------------------------------------------------ derived quantities
    def _sort_topics(self) -> None:
        """Topics in decreasing size (alpha), as after creation in the submodule (manageModels.py:540-563)."""
        order = np.argsort(-self._alphas, kind="stable")
        self._alphas = self._alphas[order]
        self._betas = self._betas[order]
        self._thetas = self._thetas.tocsc()[:, order].tocsr()
        self._tpc_labels = [self._tpc_labels[i] for i in order]

    def _derive(self) -> None:
        b = np.array(self._betas, dtype=np.float64)
        K, V = b.shape
        bs = b + 1e-12 if b.min() < 1e-12 else b
        # down-scored betas: words that are probable in every topic are pushed down (manageModels.py:594-607)
        logb = np.log(bs)
        self._betas_ds = bs * (logb - logb.mean(0, keepdims=True))
        self._topic_entropy = -(bs * logb).sum(1) / np.log(V)
        self._ndocs_active = np.asarray((self._thetas != 0).sum(0)).ravel()
        self._tpc_descriptions = [", ".join(w) for w in self.get_tpc_word_descriptions()]
        if self._ref_bow is not None:
            self._topic_coherence = np.array([metrics.npmi_coherence(b[k:k + 1], self._ref_bow) for k in range(K)])
        else:
            self._topic_coherence = np.full(K, np.nan)

    def get_tpc_word_descriptions(self, n_words: int = 15, tpc: Optional[Sequence[int]] = None) -> List[List[str]]:
        """Most relevant words per topic according to the down-scored betas."""
        tpc = range(self._betas.shape[0]) if tpc is None else tpc
        out = []
        for k in tpc:
            top = np.argsort(-self._betas_ds[k])[:n_words]
            out.append([self._vocab[i] if i < len(self._vocab) else str(i) for i in top])
        return out

    # ------------------------------------------------------------------ curation (manageModels.py:993-1244)
    def _after_edit(self, tag: str) -> int:
        self._alphas = np.asarray(self._thetas.mean(0)).ravel()
        self._derive()
        self._edits.append(tag)
        self._save_all()
        return 1

    def setTpcLabels(self, labels: Sequence[str]) -> int:
        if len(labels) != self._betas.shape[0]:
            return 0
        self._tpc_labels = [str(x) for x in labels]
        self._save_all()
        return 1

    def deleteTopics(self, tpcs: Sequence[int]) -> int:
        """Drops topics; document vectors are L1-renormalised over the remaining ones."""
        K = self._betas.shape[0]
        keep = [k for k in range(K) if k not in set(tpcs)]
        if not keep or len(keep) == K:
            return 0
        th = self._thetas.tocsc()[:, keep].tocsr().astype(np.float64)
        s = np.asarray(th.sum(1)).ravel()
        s[s == 0] = 1.0
        self._thetas = sp.diags(1.0 / s) @ th
        self._betas = self._betas[keep]
        self._tpc_labels = [self._tpc_labels[i] for i in keep]
        return self._after_edit("d " + " ".join(str(k) for k in tpcs))

    def fuseTopics(self, tpcs: Sequence[int]) -> int:
        """Merges topics into the first of them: word vector = size-weighted mean, document weight = sum."""
        tpcs = sorted(set(int(t) for t in tpcs))
        K = self._betas.shape[0]
        if len(tpcs) < 2 or tpcs[-1] >= K:
            return 0
        w = self._alphas[tpcs]
        fused = (w[None, :] @ self._betas[tpcs]) / max(w.sum(), 1e-300)
        th = self._thetas.toarray()
        th[:, tpcs[0]] = th[:, tpcs].sum(1)
        keep = [k for k in range(K) if k not in tpcs[1:]]
        self._betas[tpcs[0]] = fused
        self._betas = self._betas[keep]
        self._thetas = sp.csr_matrix(th[:, keep])
        self._tpc_labels = [self._tpc_labels[i] for i in keep]
        return self._after_edit("f " + " ".join(str(k) for k in tpcs))

    def sortTopics(self) -> int:
        self._sort_topics()
        return self._after_edit("s")

    def resetTM(self) -> int:
        """Back to the matrices the model was created with."""
        f = self._TMfolder
        self._betas = np.load(f / "betas_orig.npy")
        self._thetas = sp.load_npz(f / "thetas_orig.npz").tocsr()
        self._alphas = np.load(f / "alphas_orig.npy")
        self._tpc_labels = [f"topic {k}" for k in range(self._betas.shape[0])]
        self._sort_topics()
        self._derive()
        self._edits = []
        self._save_all()
        return 1

    def getSimilarTopics(self, npairs: int = 5):
        """Most similar topic pairs by co-occurrence in documents and by word distribution (Hellinger affinity)."""
        K = self._betas.shape[0]
        th = self._thetas.toarray()
        co = np.sqrt(th).T @ np.sqrt(th)
        wd = np.sqrt(self._betas) @ np.sqrt(self._betas).T
        iu = np.triu_indices(K, 1)
        top = lambda m: [(int(iu[0][i]), int(iu[1][i]), float(m[iu][i])) for i in np.argsort(-m[iu])[:npairs]]  # noqa: E731
        return {"Coocurring": top(co), "Worddesc": top(wd)}

    # ------------------------------------------------------------------ accessors
    def get_alphas(self):
        return self._alphas

    def get_betas(self):
        return self._betas

    def get_thetas(self):
        return self._thetas

    def showTopics(self):
        return [(k, d) for k, d in enumerate(self._tpc_descriptions)]

    def to_dataframe(self):
        import pandas as pd

        return pd.DataFrame({"alphas": self._alphas, "topic_entropy": self._topic_entropy,
                             "topic_coherence": self._topic_coherence, "ndocs_active": self._ndocs_active,
                             "tpc_descriptions": self._tpc_descriptions, "tpc_labels": self._tpc_labels})

    # rank-biased-overlap diversity and topic diversity of the current topics (manageModels.py:757-812)
    def calculate_rbo(self, k: int = 10) -> float:
        return metrics.inverted_rbo(self._betas, k)

    def calculate_topic_diversity(self, k: int = 25) -> float:
        return metrics.topic_diversity(self._betas, k)
