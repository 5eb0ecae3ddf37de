#!/usr/bin/env python
"""DSS / TSS simulation: centralized vs non-collaborative vs random-topics baseline on synthetic corpora.

Ref: experiments/dss_tss/run_simulation.py:77-739 of the reference.  For every value of the swept
parameter (number of shared "frozen" topics, or the topic-word concentration eta) and every iteration:
generate a federated synthetic corpus (LDA or ProdLDA generator), train one ProdLDA on the union of the
nodes (centralized) and one per node (non-collaborative), and score each against the ground truth with

    TSS = sum_j max_k sum_v sqrt(beta_kv beta*_jv)          (higher is better, max = n_topics)
    DSS = sum |sqrt(th*) sqrt(th*)^T - sqrt(th) sqrt(th)^T| / n_docs   (lower is better)

plus a random-topics baseline (Dirichlet draws).  Results go to a pickled pandas DataFrame with the
reference's column layout.  Training uses gfedntm_b200.models.AVITM (fused kernels on a GPU)."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np
import pandas as pd
import scipy.sparse as sp

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))

from gfedntm_b200.models import AVITM, BOWDataset  # noqa: E402
from gfedntm_b200.utils.metrics import dss, tss  # noqa: E402
from gfedntm_b200.utils.synthetic import generate_federated_corpus  # noqa: E402


def train_avitm(counts: sp.csr_matrix, n_topics: int, epochs: int, seed: int, device=None):
    vocab = [f"wd{i}" for i in range(counts.shape[1])]
    ds = BOWDataset(counts, vocab)
    m = AVITM(input_size=counts.shape[1], n_components=n_topics, model_type="prodLDA", hidden_sizes=(50, 50),
              batch_size=64, num_epochs=epochs, verbose=False, seed=seed, device=device)
    m.fit(ds)
    return m, ds


def evaluate(model, ds, topics_gt, thetas_gt, n_inf: int):
    betas = model.get_topic_word_distribution()
    thetas = model.get_doc_topic_distribution(ds, n_samples=5)
    k = min(n_inf, len(thetas))
    return tss(betas, topics_gt), dss(thetas[:k], thetas_gt[:k])


def run(cfg: dict, out_path: str, epochs: int, iters: int, device=None, scale: float = 1.0) -> pd.DataFrame:
    rows = []
    rng = np.random.default_rng(0)
    n_docs = max(int(cfg["n_docs"] * scale), 50)
    for value in cfg["values"]:
        for it in range(iters):
            beta = value if cfg["experiment"] == "eta" else cfg["beta"]
            frozen = cfg["frozen_topics"] if cfg["experiment"] == "eta" else int(value)
            topics, doc_topics, counts = generate_federated_corpus(
                vocab_size=cfg["vocab_size"], n_topics=cfg["n_topics"], beta=beta, alpha=cfg["alpha"],
                n_docs=n_docs, nwords=tuple(cfg["nwords"]), n_nodes=cfg["n_nodes"], frozen_topics=frozen,
                own_topics=cfg.get("own_topics"), seed=1000 * it + int(1e6 * float(value)) % 997)
            t0 = time.perf_counter()
            # centralized
            all_counts = sp.vstack(counts).tocsr()
            all_thetas = np.vstack(doc_topics)
            m, ds = train_avitm(all_counts, cfg["n_topics"], epochs, seed=it, device=device)
            tss_c, dss_c = evaluate(m, ds, topics, all_thetas, cfg["n_docs_global_inf"])
            # non-collaborative (mean over nodes)
            res = []
            for node in range(cfg["n_nodes"]):
                mn, dn = train_avitm(counts[node], cfg["n_topics"], epochs, seed=100 + it, device=device)
                res.append(evaluate(mn, dn, topics, doc_topics[node], cfg["n_docs_inf"]))
            tss_n, dss_n = np.mean([r[0] for r in res]), np.mean([r[1] for r in res])
            # random baseline
            rand_b = rng.dirichlet(np.full(cfg["vocab_size"], beta), cfg["n_topics"])
            rand_t = rng.dirichlet(np.full(cfg["n_topics"], cfg["alpha"]), cfg["n_docs_global_inf"])
            tss_b = tss(rand_b, topics)
            dss_b = dss(rand_t, all_thetas[:cfg["n_docs_global_inf"]])
            rows.append({cfg["experiment"]: value, "iter": it, "TSS_centralized": tss_c, "DSS_centralized": dss_c,
                         "TSS_non_colab": tss_n, "DSS_non_colab": dss_n, "TSS_baseline": tss_b,
                         "DSS_baseline": dss_b, "seconds": time.perf_counter() - t0})
            print(rows[-1], flush=True)
    df = pd.DataFrame(rows)
    os.makedirs(os.path.dirname(out_path) or ".", exist_ok=True)
    df.to_pickle(out_path)
    return df


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--config", default=os.path.join(os.path.dirname(__file__), "config/eta_variable/config.json"))
    ap.add_argument("--out", default=None)
    ap.add_argument("--epochs", type=int, default=100)
    ap.add_argument("--iters", type=int, default=None)
    ap.add_argument("--scale", type=float, default=1.0, help="scale the documents per node (quick runs)")
    ap.add_argument("--device", default=None)
    a = ap.parse_args(argv)
    with open(a.config) as f:
        cfg = json.load(f)
    out = a.out or os.path.join(os.path.dirname(a.config).replace("config", "results"), "results.pickle")
    df = run(cfg, out, a.epochs, a.iters or cfg["iters"], device=a.device, scale=a.scale)
    print(df.groupby(cfg["experiment"]).mean(numeric_only=True))


if __name__ == "__main__":
    main()
