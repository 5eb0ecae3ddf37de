#!/usr/bin/env python
"""Word-mover's distance between the topic descriptions of node models and a centralized model.
Ref: aux_scripts/evaluation/wmd.py:13-110 (gensim ``word2vec-google-news-300``).  gensim and the
pretrained vectors are not available offline on the target boxes, so the embedding source is a
parameter: any ``{word: vector}`` mapping (e.g. loaded from a .npz / word2vec text file).  The distance
itself is computed here: the exact word-mover's distance (optimal transport between the two normalised
bags of words under the Euclidean ground cost, the quantity gensim's ``wmdistance`` returns), solved as a small
linear programme with scipy's HiGHS; the relaxed nearest-neighbour bound is kept as a cheap lower bound."""
import argparse
import numpy as np


def wmd(words_a, words_b, vectors) -> float:
    """Exact word-mover's distance between two word lists (uniform nBOW weights over the in-vocabulary words,
    repeated words pooled)."""
    from scipy.optimize import linprog

    def bag(ws):
        cnt = {}
        for w in ws:
            if w in vectors:
                cnt[w] = cnt.get(w, 0) + 1
        keys = sorted(cnt)
        wts = np.array([cnt[k] for k in keys], dtype=np.float64)
        return keys, (wts / wts.sum() if len(keys) else wts)

    ka, pa = bag(words_a)
    kb, pb = bag(words_b)
    if not ka or not kb:
        return float("nan")
    A = np.array([vectors[w] for w in ka], dtype=np.float64)
    Bm = np.array([vectors[w] for w in kb], dtype=np.float64)
    cost = np.linalg.norm(A[:, None, :] - Bm[None, :, :], axis=2)
    na, nb = len(ka), len(kb)
    # transport plan T >= 0 with row sums pa and column sums pb, minimise <T, cost>
    Aeq = np.zeros((na + nb, na * nb))
    for i in range(na):
        Aeq[i, i * nb:(i + 1) * nb] = 1.0
    for j in range(nb):
        Aeq[na + j, j::nb] = 1.0
    res = linprog(cost.ravel(), A_eq=Aeq[:-1], b_eq=np.concatenate([pa, pb])[:-1], bounds=(0, None), method="highs")
    return float(res.fun) if res.success else float("nan")


def relaxed_wmd(words_a, words_b, vectors) -> float:
    a = np.array([vectors[w] for w in words_a if w in vectors])
    b = np.array([vectors[w] for w in words_b if w in vectors])
    if len(a) == 0 or len(b) == 0:
        return float("nan")
    d = np.linalg.norm(a[:, None, :] - b[None, :, :], axis=2)
    return float(max(d.min(1).mean(), d.min(0).mean()))


def topic_distance_matrix(topics_a, topics_b, vectors, exact: bool = True) -> np.ndarray:
    fn = wmd if exact else relaxed_wmd
    return np.array([[fn(ta, tb, vectors) for tb in topics_b] for ta in topics_a])


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--node_model", required=True, help="npz written by a client (key `topics`)")
    ap.add_argument("--central_model", required=True, help="npz of the centralized / global model with `topics`")
    ap.add_argument("--vectors", required=True, help="npz with arrays `words` and `vectors`")
    ap.add_argument("--relaxed", action="store_true", help="nearest-neighbour lower bound instead of the exact distance")
    a = ap.parse_args(argv)
    vz = np.load(a.vectors, allow_pickle=True)
    vectors = dict(zip(vz["words"].tolist(), vz["vectors"]))
    ta = np.load(a.node_model, allow_pickle=True)["topics"].tolist()
    tb = np.load(a.central_model, allow_pickle=True)["topics"].tolist()
    m = topic_distance_matrix(ta, tb, vectors, exact=not a.relaxed)
    print("mean best-match WMD:", np.nanmean(np.nanmin(m, axis=1)))


if __name__ == "__main__":
    main()
