"""Topic-model quality metrics that need no external corpus tooling.

* TSS / DSS against a synthetic ground truth — ref experiments/dss_tss/run_simulation.py:321-355 and
  federated_avitm.py:152-193.
* NPMI coherence, topic diversity, inverted RBO — native replacements for the gensim / rbo based
  ``TMmodel`` metrics of the vendored topicmodeler (manageModels.py:649-812), which are not installed
  on the target boxes.
"""
from __future__ import annotations

from typing import Sequence

import numpy as np
import scipy.sparse as sp


def tss(betas: np.ndarray, betas_gt: np.ndarray) -> float:
    """Topic similarity score: sum over true topics of the best Bhattacharyya match,
    ``sum_j max_k sum_v sqrt(beta_kv * beta*_jv)`` (max = number of true topics)."""
    sim = np.sqrt(np.asarray(betas, dtype=np.float64)) @ np.sqrt(np.asarray(betas_gt, dtype=np.float64)).T
    return float(np.sum(np.max(sim, axis=0)))


def dss(thetas: np.ndarray, thetas_gt: np.ndarray, block: int = 2048) -> float:
    """Document similarity score: ``sum |sqrt(th*) sqrt(th*)^T - sqrt(th) sqrt(th)^T| / n_docs``
    (lower is better), evaluated in row blocks so the n_docs^2 matrices are never held at once."""
    a = np.sqrt(np.asarray(thetas_gt, dtype=np.float64))
    b = np.sqrt(np.asarray(thetas, dtype=np.float64))
    n = a.shape[0]
    tot = 0.0
    for s in range(0, n, block):
        tot += np.abs(a[s:s + block] @ a.T - b[s:s + block] @ b.T).sum()
    return float(tot / n)


def top_words(betas: np.ndarray, k: int = 10) -> np.ndarray:
    return np.argsort(-np.asarray(betas), axis=1)[:, :k]


def topic_diversity(betas: np.ndarray, k: int = 25) -> float:
    """Fraction of unique words among the top-``k`` words of all topics."""
    tw = top_words(betas, k)
    return float(len(np.unique(tw)) / tw.size)


def npmi_coherence(betas: np.ndarray, bow: sp.spmatrix, k: int = 10, eps: float = 1e-12) -> float:
    """Mean NPMI over the top-``k`` word pairs of each topic with document co-occurrence counts taken
    from ``bow`` (documents x vocabulary).  Values in [-1, 1], higher is better."""
    X = sp.csr_matrix(bow)
    X = (X > 0).astype(np.float64)
    n_docs = X.shape[0]
    tw = top_words(betas, k)
    words = np.unique(tw)
    sub = X[:, words].tocsc()
    df = np.asarray(sub.sum(axis=0)).ravel() / n_docs
    co = (sub.T @ sub).toarray() / n_docs
    pos = {w: i for i, w in enumerate(words)}
    scores = []
    for t in range(tw.shape[0]):
        vals = []
        for i in range(k):
            for j in range(i + 1, k):
                a, b = pos[tw[t, i]], pos[tw[t, j]]
                pij = co[a, b]
                if pij <= 0:
                    vals.append(-1.0)
                    continue
                pmi = np.log(pij / (df[a] * df[b] + eps))
                vals.append(pmi / (-np.log(pij + eps)))
        scores.append(np.mean(vals))
    return float(np.mean(scores))


def _rbo(l1: Sequence[int], l2: Sequence[int], p: float = 0.9) -> float:
    """Rank-biased overlap (extrapolated) of two equally long rankings."""
    k = min(len(l1), len(l2))
    s1, s2 = set(), set()
    overlap, summ = 0, 0.0
    for d in range(1, k + 1):
        a, b = l1[d - 1], l2[d - 1]
        if a == b:
            overlap += 1
        else:
            overlap += (a in s2) + (b in s1)
        s1.add(a)
        s2.add(b)
        summ += overlap / d * p ** d
    return float(overlap / k * p ** k + (1 - p) / p * summ)


def inverted_rbo(betas: np.ndarray, k: int = 10, p: float = 0.9) -> float:
    """1 - mean pairwise RBO of the topics' top-``k`` word rankings (higher = more diverse)."""
    tw = top_words(betas, k)
    n = tw.shape[0]
    vals = [_rbo(list(tw[i]), list(tw[j]), p) for i in range(n) for j in range(i + 1, n)]
    return float(1.0 - np.mean(vals)) if vals else 1.0


def thresholded_thetas(thetas: np.ndarray, thr: float = 3e-3) -> np.ndarray:
    """Zero entries below ``thr`` and L1-renormalise rows (federated_model.py:170-173)."""
    t = np.array(thetas, dtype=np.float64, copy=True)
    t[t < thr] = 0
    s = t.sum(axis=1, keepdims=True)
    s[s == 0] = 1.0
    return t / s
