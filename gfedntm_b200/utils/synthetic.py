"""Synthetic LDA corpora with known ground truth.

Ref: src/utils/generate_synthetic.py:34-96 (federated layout: per-node corpora with `frozen`
shared topics and node-specific topics; tokens ``wd<idx>``; npz keys ``documents``, ``vocab_size``,
``topic_vectors``, ``doc_topics``) and experiments/dss_tss/run_simulation.py:166-174 (ProdLDA-style
generator).  Generation is vectorised (multinomial draws per document) and can run on the GPU for
large vocabularies.
"""
from __future__ import annotations

from typing import List, Optional, Tuple

import numpy as np
import scipy.sparse as sp


def rotate_array(arr: np.ndarray, k: int) -> np.ndarray:
    """Cyclic right rotation by ``k`` (the reference's ``rotateArray``)."""
    k = k % len(arr)
    return np.concatenate((arr[-k:], arr[:-k])) if k else arr.copy()


def generate_topics(vocab_size: int, n_topics: int, eta: float, rng: np.random.Generator) -> np.ndarray:
    return rng.dirichlet(np.full(vocab_size, eta), n_topics)


def node_prior(n_topics: int, frozen_topics: int, own_topics: int, node: int, alpha: Optional[float] = None
               ) -> np.ndarray:
    """Dirichlet prior of one node: ``frozen_topics`` shared topics and ``own_topics`` node topics at
    ``alpha``, the rest at ``alpha/10000``; the non-shared part is rotated per node."""
    alpha = 1.0 / n_topics if alpha is None else alpha
    rest = n_topics - frozen_topics
    own = np.array([alpha] * min(own_topics, rest) + [alpha / 10000] * max(rest - own_topics, 0))
    own = rotate_array(own, node * own_topics) if rest else own
    return np.concatenate((np.full(frozen_topics, alpha), own))


def sample_corpus_counts(theta: np.ndarray, topics: np.ndarray, doc_lens: np.ndarray,
                         rng: np.random.Generator, prod: bool = False) -> sp.csr_matrix:
    """Bag-of-words counts for documents with topic mixtures ``theta`` [N,K].

    LDA generator: word ~ Cat(theta . beta).  ProdLDA generator (``prod=True``): word ~
    Cat(softmax-like product of experts  prod_k beta_k^{theta_k}).
    """
    N = theta.shape[0]
    rows, cols, vals = [], [], []
    if prod:
        logb = np.log(topics + 1e-30)
    for d in range(N):
        if prod:
            lp = theta[d] @ logb
            p = np.exp(lp - lp.max())
            p /= p.sum()
        else:
            p = theta[d] @ topics
            p /= p.sum()
        c = rng.multinomial(int(doc_lens[d]), p)
        nz = np.nonzero(c)[0]
        rows.append(np.full(len(nz), d))
        cols.append(nz)
        vals.append(c[nz])
    m = sp.csr_matrix((np.concatenate(vals).astype(np.float32),
                       (np.concatenate(rows), np.concatenate(cols))), shape=(N, topics.shape[1]))
    m.indptr = m.indptr.astype(np.int32)
    m.indices = m.indices.astype(np.int32)
    return m


def counts_to_documents(counts: sp.csr_matrix) -> List[List[str]]:
    """CSR counts -> token lists ``['wd12', 'wd12', 'wd7', ...]`` (token ids start at 0 like the
    reference generator)."""
    docs = []
    for d in range(counts.shape[0]):
        s, e = counts.indptr[d], counts.indptr[d + 1]
        toks = np.repeat(counts.indices[s:e], counts.data[s:e].astype(int))
        docs.append(["wd" + str(t) for t in toks])
    return docs


def generate_federated_corpus(vocab_size: int = 5000, n_topics: int = 50, beta: float = 1e-2,
                              alpha: Optional[float] = None, n_docs: int = 1000, nwords=(150, 250),
                              n_nodes: int = 5, frozen_topics: int = 5, own_topics: Optional[int] = None,
                              seed: int = 0, prod: bool = False):
    """Per-node synthetic corpora.  Returns ``(topic_vectors [K,V], doc_topics [nodes][N,K],
    counts [nodes] csr)``.  Defaults are the reference script's constants."""
    rng = np.random.default_rng(seed)
    if own_topics is None:
        own_topics = int((n_topics - frozen_topics) / n_nodes)
    topics = generate_topics(vocab_size, n_topics, beta, rng)
    doc_topics, counts = [], []
    for node in range(n_nodes):
        prior = node_prior(n_topics, frozen_topics, own_topics, node, alpha)
        theta = rng.dirichlet(prior, n_docs)
        lens = rng.integers(nwords[0], nwords[1], n_docs)
        doc_topics.append(theta)
        counts.append(sample_corpus_counts(theta, topics, lens, rng, prod=prod))
    return topics, doc_topics, counts


def save_synthetic_npz(path: str, topics: np.ndarray, doc_topics, counts) -> None:
    """``synthetic.npz`` in the layout main.py expects (main.py:138-146)."""
    documents = np.empty(len(counts), dtype=object)
    for i, c in enumerate(counts):
        documents[i] = counts_to_documents(c)
    np.savez(path, vocab_size=topics.shape[1], n_topics=topics.shape[0], frozen_topics=0,
             beta=0.0, alpha=0.0, n_docs=counts[0].shape[0], topic_vectors=topics,
             doc_topics=np.asarray(doc_topics), documents=documents)


def generate_corpus_gpu(n_docs: int, vocab_size: int, n_topics: int, eta: float = 0.01,
                        nwords=(150, 250), seed: int = 0, device="cuda") -> Tuple[sp.csr_matrix, np.ndarray]:
    """Large-vocabulary LDA corpus sampled with torch (per-topic ``torch.multinomial`` draws), for
    benchmark shapes such as V=200k / K=200 where the host loop above would take minutes."""
    import torch

    g = torch.Generator(device=device).manual_seed(seed)
    dev = torch.device(device)
    conc = torch.full((vocab_size,), eta, device=dev)
    topics = torch._sample_dirichlet(conc.expand(n_topics, vocab_size).contiguous(), generator=g)
    topics = topics.clamp_min(1e-30)
    topics = topics / topics.sum(1, keepdim=True)
    alpha = torch.full((n_docs, n_topics), 1.0 / n_topics, device=dev)
    theta = torch._sample_dirichlet(alpha, generator=g).clamp_min(1e-30)
    theta = theta / theta.sum(1, keepdim=True)
    lens = torch.randint(nwords[0], nwords[1], (n_docs,), device=dev, generator=g)
    L = int(lens.max().item())
    z = torch.multinomial(theta, L, replacement=True, generator=g)           # [N, L] topic of each token
    valid = torch.arange(L, device=dev)[None, :] < lens[:, None]
    words = torch.zeros(n_docs, L, dtype=torch.long, device=dev)
    zf = z[valid]
    wf = torch.zeros_like(zf)
    order = torch.argsort(zf)
    counts = torch.bincount(zf, minlength=n_topics).tolist()
    pos = 0
    for k, n in enumerate(counts):
        if n:
            wf[order[pos:pos + n]] = torch.multinomial(topics[k], n, replacement=True, generator=g)
            pos += n
    words[valid] = wf
    doc_id = torch.arange(n_docs, device=dev)[:, None].expand(n_docs, L)[valid]
    key = doc_id * vocab_size + wf
    uniq, cnt = torch.unique(key, return_counts=True)
    rows = (uniq // vocab_size).cpu().numpy()
    cols = (uniq % vocab_size).cpu().numpy()
    m = sp.csr_matrix((cnt.float().cpu().numpy(), (rows, cols)), shape=(n_docs, vocab_size))
    m.indptr = m.indptr.astype(np.int32)
    m.indices = m.indices.astype(np.int32)
    return m, topics.cpu().numpy()
