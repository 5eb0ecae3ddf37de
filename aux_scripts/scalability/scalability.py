#!/usr/bin/env python
"""Scalability grid: train time / resource usage of the trainers over corpus sizes and topic counts.

Role of the submodule's ``aux_scripts/scalability*.py`` (grid over trainers with the ``Mem`` sampler running
and the coherence timed separately).  For every (trainer, n_docs, n_topics) cell a synthetic LDA corpus of
that size is generated, the model is trained through ``TMWrapper`` while ``ResourceSampler`` polls the
process, and one CSV row is written: train seconds, docs/s, peak RSS, peak GPU memory (when NVML is there),
NPMI coherence and the seconds it took to compute it.

    python aux_scripts/scalability/scalability.py --docs 2000 8000 --topics 10 50 --epochs 5 --out scal.csv
"""
import argparse
import csv
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def peak_from_samples(path: str):
    rss, gpu = 0, 0.0
    with open(path) as f:
        rd = csv.DictReader(f)
        for row in rd:
            rss = max(rss, int(row["rss"]))
            for k, v in row.items():
                if k.endswith("_mem_used") and v:
                    gpu = max(gpu, float(v))
    return rss, gpu


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--trainers", nargs="+", default=["prod"], choices=["prod", "lda"])
    ap.add_argument("--docs", nargs="+", type=int, default=[2000, 8000])
    ap.add_argument("--topics", nargs="+", type=int, default=[10, 50])
    ap.add_argument("--vocab", type=int, default=2000)
    ap.add_argument("--epochs", type=int, default=5)
    ap.add_argument("--device", default=None)
    ap.add_argument("--out", default="scalability.csv")
    a = ap.parse_args(argv)

    from gfedntm_b200.aux_modules.tm_wrapper import TMWrapper
    from gfedntm_b200.utils import metrics
    from gfedntm_b200.utils.mem_usage import ResourceSampler
    from gfedntm_b200.utils.synthetic import counts_to_documents, generate_topics, sample_corpus_counts

    tw = TMWrapper()
    rows = []
    rng = np.random.default_rng(0)
    kw = {"device": a.device} if a.device else {}
    for n_docs in a.docs:
        topics = generate_topics(a.vocab, max(a.topics), 0.02, rng)
        theta = rng.dirichlet(np.full(topics.shape[0], 1.0 / topics.shape[0]), n_docs)
        counts = sample_corpus_counts(theta, topics, rng.integers(80, 120, n_docs), rng).tocsr()
        corpus = [" ".join(d) for d in counts_to_documents(counts)]
        for trainer in a.trainers:
            for k in a.topics:
                with tempfile.TemporaryDirectory() as td:
                    usage = os.path.join(td, "usage.csv")
                    with ResourceSampler(usage, period_s=0.2, gpu=True):
                        model, ds, secs = tw.train_root_model(corpus, trainer=trainer, ntopics=k,
                                                              num_epochs=a.epochs, max_features=None, **kw)
                    rss, gpu = peak_from_samples(usage)
                t0 = time.perf_counter()
                cohr = metrics.npmi_coherence(model.get_topic_word_distribution(), ds.X)
                rows.append({"trainer": trainer, "n_docs": n_docs, "n_topics": k, "epochs": a.epochs,
                             "backend": model.backend, "train_s": round(secs, 3),
                             "docs_per_s": round(n_docs * a.epochs / max(secs, 1e-9), 1), "peak_rss": rss,
                             "peak_gpu_mem": gpu, "npmi": round(float(cohr), 4),
                             "coherence_s": round(time.perf_counter() - t0, 3)})
                print(rows[-1], flush=True)
    with open(a.out, "w", newline="") as f:
        w = csv.DictWriter(f, fieldnames=list(rows[0].keys()))
        w.writeheader()
        w.writerows(rows)
    return a.out


if __name__ == "__main__":
    main()
