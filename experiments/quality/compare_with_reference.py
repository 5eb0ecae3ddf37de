#!/usr/bin/env python
"""Topic quality of the fused sm_100a trainer against the unmodified reference on ONE synthetic corpus.

Both arms train ProdLDA with the same hyper-parameters on the same LDA-generated corpus (generator of
src/utils/generate_synthetic.py:34-79: topics ~ Dir(eta), theta ~ Dir(alpha), ancestral sampling) and are
scored with the reference's own measures: TSS against the generating topics
(experiments/dss_tss/run_simulation.py:321-334), NPMI coherence and topic diversity.  The reference arm
imports the stock sources from ``baseline/_ref`` (see baseline/install_ref.sh).

    python experiments/quality/compare_with_reference.py --epochs 30 --out profiles/quality_r1.json
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--docs", type=int, default=8000)
    ap.add_argument("--vocab", type=int, default=2000)
    ap.add_argument("--topics", type=int, default=20)
    ap.add_argument("--eta", type=float, default=0.02)
    ap.add_argument("--epochs", type=int, default=30)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--seeds", type=int, default=2)
    ap.add_argument("--device", default="cuda")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()

    import torch

    from gfedntm_b200.models import AVITM, BOWDataset
    from gfedntm_b200.utils import metrics
    from gfedntm_b200.utils.synthetic import generate_topics, sample_corpus_counts

    rng = np.random.default_rng(7)
    topics = generate_topics(a.vocab, a.topics, a.eta, rng)
    theta = rng.dirichlet(np.full(a.topics, 1.0 / a.topics), a.docs)
    lens = rng.integers(150, 250, a.docs)
    counts = sample_corpus_counts(theta, topics, lens, rng).tocsr().astype(np.float32)
    idx2token = {i: f"wd{i}" for i in range(a.vocab)}
    hp = dict(input_size=a.vocab, n_components=a.topics, model_type="prodLDA", hidden_sizes=(50, 50),
              activation="softplus", dropout=0.2, learn_priors=True, batch_size=a.batch, lr=2e-3, momentum=0.99,
              solver="adam", num_epochs=a.epochs, reduce_on_plateau=False, num_data_loader_workers=0, verbose=False)

    def score(betas):
        betas = np.asarray(betas, dtype=np.float64)
        return {"tss": metrics.tss(betas, topics), "npmi": metrics.npmi_coherence(betas, counts),
                "diversity": metrics.topic_diversity(betas)}

    rows = []
    rand = rng.dirichlet(np.full(a.vocab, a.eta), a.topics)
    rows.append(dict(arm="random topics", seed=-1, seconds=0.0, **score(rand)))
    rows.append(dict(arm="generating topics", seed=-1, seconds=0.0, **score(topics)))

    for seed in range(a.seeds):
        torch.manual_seed(seed)
        m = AVITM(device=a.device, seed=seed, **hp)
        t0 = time.perf_counter()
        m.fit(BOWDataset(counts, idx2token))
        if a.device.startswith("cuda"):
            torch.cuda.synchronize()
        rows.append(dict(arm=f"ours ({m.backend})", seed=seed, seconds=time.perf_counter() - t0,
                         **score(m.get_topic_word_distribution())))

    try:
        from baseline.ref_arm import _import_reference
        import logging

        RefAVITM, _, RefBOW, *_ = _import_reference()
        dense = np.asarray(counts.todense(), dtype=np.float32)
        for seed in range(a.seeds):
            torch.manual_seed(seed)
            np.random.seed(seed)
            ref_hp = dict(hp)
            ref_hp.update(topic_prior_mean=0.0, topic_prior_variance=None, num_samples=20)
            rm = RefAVITM(logger=logging.getLogger("ref"), **ref_hp)
            t0 = time.perf_counter()
            rm.fit(RefBOW(dense, idx2token))
            if torch.cuda.is_available():
                torch.cuda.synchronize()
            rows.append(dict(arm="reference (stock PyTorch)", seed=seed, seconds=time.perf_counter() - t0,
                             **score(rm.get_topic_word_distribution())))
    except Exception as ex:  # pragma: no cover - the reference arm is optional
        rows.append(dict(arm=f"reference unavailable: {type(ex).__name__}: {ex}"[:160], seed=-1, seconds=0.0,
                         tss=float("nan"), npmi=float("nan"), diversity=float("nan")))

    print(f"{'arm':34s} seed   TSS(max {a.topics})   NPMI   diversity   train s")
    for r in rows:
        print(f"{r['arm']:34s} {r['seed']:4d}   {r['tss']:8.3f}    {r['npmi']:6.3f}   {r['diversity']:6.3f}   {r['seconds']:8.1f}")
    if a.out:
        with open(a.out, "w") as f:
            json.dump({"config": vars(a), "rows": rows}, f, indent=1)


if __name__ == "__main__":
    main()
