#!/usr/bin/env python
"""Word-mover's distance between the topic descriptions of node models and a centralized model.
Ref: aux_scripts/evaluation/wmd.py:13-110 (gensim ``word2vec-google-news-300``).  gensim and the
pretrained vectors are not available offline on the target boxes, so the embedding source is a
parameter: any ``{word: vector}`` mapping (e.g. loaded from a .npz / word2vec text file).  The distance
itself is computed here (relaxed WMD: mean nearest-neighbour cost in both directions)."""
import argparse
import numpy as np


def relaxed_wmd(words_a, words_b, vectors) -> float:
    a = np.array([vectors[w] for w in words_a if w in vectors])
    b = np.array([vectors[w] for w in words_b if w in vectors])
    if len(a) == 0 or len(b) == 0:
        return float("nan")
    d = np.linalg.norm(a[:, None, :] - b[None, :, :], axis=2)
    return float(max(d.min(1).mean(), d.min(0).mean()))


def topic_distance_matrix(topics_a, topics_b, vectors) -> np.ndarray:
    return np.array([[relaxed_wmd(ta, tb, vectors) for tb in topics_b] for ta in topics_a])


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--node_model", required=True, help="npz written by a client (key `topics`)")
    ap.add_argument("--central_model", required=True, help="npz of the centralized / global model with `topics`")
    ap.add_argument("--vectors", required=True, help="npz with arrays `words` and `vectors`")
    a = ap.parse_args(argv)
    vz = np.load(a.vectors, allow_pickle=True)
    vectors = dict(zip(vz["words"].tolist(), vz["vectors"]))
    ta = np.load(a.node_model, allow_pickle=True)["topics"].tolist()
    tb = np.load(a.central_model, allow_pickle=True)["topics"].tolist()
    m = topic_distance_matrix(ta, tb, vectors)
    print("mean best-match WMD:", np.nanmean(np.nanmin(m, axis=1)))


if __name__ == "__main__":
    main()
