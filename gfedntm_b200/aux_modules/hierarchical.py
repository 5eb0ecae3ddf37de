"""Second-level (hierarchical) corpus construction: HTM-WS and HTM-DS.

The submodule's ``HierarchicalTMManager`` (src/aux_modules/topicmodeler/src/topicmodeling/topicmodeling.py:
1562-1830) expands one topic ``t`` of a trained first-level model into a sub-corpus on which a child model
is trained:

* **HTM-DS** (document selection): keep the documents whose weight of topic ``t`` is at least ``thr``.
* **HTM-WS** (word selection): keep, in every document, the word occurrences assigned to topic ``t``, where
  the assignment of word ``w`` in document ``d`` follows ``p(z | w, d) ~ theta[d, z] * beta[z, w]``; the
  assignment is either the arg-max or a draw from that distribution.

Here both operate on the sparse count matrix directly (no token strings round trip) and are vectorised over
the non-zeros, so they also run on large vocabularies.
"""
from __future__ import annotations

from typing import Optional, Tuple

import numpy as np
import scipy.sparse as sp


def htm_ds(counts: sp.spmatrix, thetas: np.ndarray, topic: int, thr: float = 0.1) -> Tuple[sp.csr_matrix, np.ndarray]:
    """Returns (sub-corpus, indices of the kept documents)."""
    counts = sp.csr_matrix(counts)
    thetas = np.asarray(thetas.todense() if sp.issparse(thetas) else thetas)
    keep = np.flatnonzero(thetas[:, topic] >= thr)
    return counts[keep], keep


def htm_ws(counts: sp.spmatrix, thetas: np.ndarray, betas: np.ndarray, topic: int, *, sample: bool = False,
           seed: Optional[int] = None, min_words: int = 1) -> Tuple[sp.csr_matrix, np.ndarray]:
    """Word-selection sub-corpus for ``topic``.  ``sample=False``: an occurrence is kept when ``topic`` is the
    arg-max of ``theta[d,:] * beta[:,w]``; ``sample=True``: every occurrence is assigned by a draw from the
    normalised product and kept when the draw equals ``topic``.  Documents left with fewer than ``min_words``
    words are dropped.  Returns (sub-corpus over the same vocabulary, indices of the kept documents)."""
    m = sp.coo_matrix(counts)
    thetas = np.asarray(thetas.todense() if sp.issparse(thetas) else thetas, dtype=np.float64)
    betas = np.asarray(betas, dtype=np.float64)
    rows, cols, vals = m.row, m.col, m.data.astype(np.int64)
    new_vals = np.zeros_like(vals)
    rng = np.random.default_rng(seed)
    step = 1 << 16
    for s in range(0, rows.size, step):
        r, c, v = rows[s:s + step], cols[s:s + step], vals[s:s + step]
        post = thetas[r] * betas[:, c].T  # [nnz_block, K]
        if not sample:
            new_vals[s:s + step] = np.where(post.argmax(1) == topic, v, 0)
        else:
            z = post.sum(1, keepdims=True)
            p_t = np.divide(post[:, topic:topic + 1], z, out=np.zeros_like(z), where=z > 0)[:, 0]
            new_vals[s:s + step] = rng.binomial(v, np.clip(p_t, 0.0, 1.0))
    sub = sp.csr_matrix((new_vals, (rows, cols)), shape=m.shape)
    sub.eliminate_zeros()
    keep = np.flatnonzero(np.asarray(sub.sum(1)).ravel() >= min_words)
    return sub[keep], keep
