#!/usr/bin/env python
"""Synthetic federated LDA corpus generator (CLI).
Ref: src/utils/generate_synthetic.py:34-96 — same constants by default and the same outputs: per-node
``node<i>.txt`` (Mallet format ``<id> 0 w w ...``), ``node<i>.npz`` and ``synthetic_all_nodes.npz`` /
``synthetic.npz`` with keys ``documents, vocab_size, topic_vectors, doc_topics, ...``."""
from __future__ import annotations

import argparse
import os

import numpy as np

from .synthetic import counts_to_documents, generate_federated_corpus, save_synthetic_npz


def main(argv=None) -> str:
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--vocab_size", type=int, default=5000)
    ap.add_argument("--n_topics", type=int, default=50)
    ap.add_argument("--beta", type=float, default=1e-2, help="Dirichlet concentration of the topic-word vectors")
    ap.add_argument("--alpha", type=float, default=None, help="doc-topic concentration (default 1/n_topics)")
    ap.add_argument("--n_docs", type=int, default=1000, help="documents per node")
    ap.add_argument("--n_nodes", type=int, default=5)
    ap.add_argument("--frozen_topics", type=int, default=5)
    ap.add_argument("--nwords", type=int, nargs=2, default=(150, 250))
    ap.add_argument("--prod", action="store_true", help="ProdLDA-style generator (product of experts)")
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--out_dir", type=str, default="static/datasets")
    a = ap.parse_args(argv)
    topics, doc_topics, counts = generate_federated_corpus(
        vocab_size=a.vocab_size, n_topics=a.n_topics, beta=a.beta, alpha=a.alpha, n_docs=a.n_docs,
        nwords=tuple(a.nwords), n_nodes=a.n_nodes, frozen_topics=a.frozen_topics, seed=a.seed, prod=a.prod)
    os.makedirs(a.out_dir, exist_ok=True)
    for i, c in enumerate(counts):
        docs = counts_to_documents(c)
        with open(os.path.join(a.out_dir, f"node{i + 1}.txt"), "w") as f:
            for d, toks in enumerate(docs):
                f.write(f"{d} 0 {' '.join(toks)}\n")
        np.savez(os.path.join(a.out_dir, f"node{i + 1}.npz"), documents=np.array(docs, dtype=object),
                 doc_topics=doc_topics[i], topic_vectors=topics, vocab_size=a.vocab_size)
    path = os.path.join(a.out_dir, "synthetic.npz")
    save_synthetic_npz(path, topics, doc_topics, counts)
    save_synthetic_npz(os.path.join(a.out_dir, "synthetic_all_nodes.npz"), topics, doc_topics, counts)
    print(f"wrote {path} ({a.n_nodes} nodes x {a.n_docs} docs, V={a.vocab_size}, K={a.n_topics})")
    return path


if __name__ == "__main__":
    main()
