#!/usr/bin/env python
"""Topic quality at the configuration the reference publishes numbers for.

Configuration (experiments/dss_tss/config/eta_variable/config.json of the reference, eta = 0.01): 5 nodes x 10 000
documents, V = 5000, K = 50, 10 shared ("frozen") topics, ProdLDA with hidden sizes (100, 100), batch 64, 100 epochs
(run_simulation.py:296-308 there).  Published (results/eta_variable/results.pickle): TSS 8.679 +- 0.042 centralized,
7.571 +- 0.048 non-collaborative, 3.564 +- 0.098 random topics; DSS 2555.5 / 3066.7 / 834.6.

Arms trained here, all on the same generated corpus:
  centralized          one fused ProdLDA on the union of the nodes
  non_collaborative    one fused ProdLDA per node (mean over nodes)
  federated_grad       5 clients, per-minibatch aggregation: weighted gradients + one global Adam (agg="grad")
  federated_param      5 clients, local Adam + weighted parameter mean (agg="param", the reference's semantics)
  reference            (--reference) the unmodified reference AVITM.fit on the union, same box
The federated arms run all clients on one GPU through the emulated-ranks aggregation kernel
(gfedntm_b200/federated/emulated.py): same kernels and arithmetic as one client per GPU.

    python experiments/quality/published_config.py --epochs 100 --out profiles/quality_published_r2.json
"""
import argparse
import json
import os
import sys
import time

import numpy as np
import scipy.sparse as sp

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--epochs", type=int, default=100)
    ap.add_argument("--docs", type=int, default=10000, help="documents per node")
    ap.add_argument("--seeds", type=int, default=1)
    ap.add_argument("--reference", action="store_true")
    ap.add_argument("--reference_epochs", type=int, default=None)
    ap.add_argument("--device", default="cuda:0")
    ap.add_argument("--out", default=None)
    a = ap.parse_args()

    import torch

    from gfedntm_b200.federated.emulated import EmulatedFederation
    from gfedntm_b200.models import AVITM, BOWDataset
    from gfedntm_b200.utils import metrics
    from gfedntm_b200.utils.synthetic import generate_federated_corpus

    cfg = json.load(open(os.path.join(ROOT, "experiments/dss_tss/config/eta_variable/config.json")))
    V, K, n_nodes = cfg["vocab_size"], cfg["n_topics"], cfg["n_nodes"]
    rows = []
    for seed in range(a.seeds):
        topics, doc_topics, counts = generate_federated_corpus(
            vocab_size=V, n_topics=K, beta=0.01, alpha=cfg["alpha"], n_docs=a.docs, nwords=tuple(cfg["nwords"]),
            n_nodes=n_nodes, frozen_topics=cfg["frozen_topics"], own_topics=cfg.get("own_topics"), seed=4242 + seed)
        all_counts = sp.vstack(counts).tocsr().astype(np.float32)
        all_thetas = np.vstack(doc_topics)
        vocab = [f"wd{i}" for i in range(V)]
        hp = dict(input_size=V, n_components=K, model_type="prodLDA", hidden_sizes=(100, 100), batch_size=64,
                  num_epochs=a.epochs, verbose=False, device=a.device)

        def score(model, ds, thetas_gt, n_inf):
            betas = np.asarray(model.get_topic_word_distribution(), dtype=np.float64)
            th = model.get_doc_topic_distribution(ds, n_samples=5)[:n_inf]
            return {"tss": metrics.tss(betas, topics), "dss": metrics.dss(th, thetas_gt[:n_inf]),
                    "npmi": metrics.npmi_coherence(betas, all_counts), "diversity": metrics.topic_diversity(betas)}

        rng = np.random.default_rng(seed)
        rand_b = rng.dirichlet(np.full(V, 0.01), K)
        rows.append(dict(arm="random_topics", seed=seed, seconds=0.0, tss=metrics.tss(rand_b, topics),
                         npmi=metrics.npmi_coherence(rand_b, all_counts), diversity=metrics.topic_diversity(rand_b)))
        # ---- centralized
        t0 = time.perf_counter()
        ds_all = BOWDataset(all_counts, vocab)
        m = AVITM(seed=seed, **hp)
        m.fit(ds_all)
        torch.cuda.synchronize()
        rows.append(dict(arm="centralized", seed=seed, seconds=time.perf_counter() - t0, backend=m.backend,
                         **score(m, ds_all, all_thetas, cfg["n_docs_global_inf"])))
        print(rows[-1], flush=True)
        # ---- non-collaborative
        t0 = time.perf_counter()
        res = []
        for node in range(n_nodes):
            dn = BOWDataset(counts[node].astype(np.float32), vocab)
            mn = AVITM(seed=100 + seed, **hp)
            mn.fit(dn)
            res.append(score(mn, dn, doc_topics[node], cfg["n_docs_inf"]))
        rows.append(dict(arm="non_collaborative", seed=seed, seconds=time.perf_counter() - t0,
                         **{k: float(np.mean([r[k] for r in res])) for k in res[0]}))
        print(rows[-1], flush=True)
        # ---- federated (5 clients, emulated ranks on one GPU)
        for agg in ("grad", "param"):
            t0 = time.perf_counter()
            trainers = [AVITM(seed=200 + seed, **hp) for _ in range(n_nodes)]
            dsets = [BOWDataset(counts[n].astype(np.float32), vocab) for n in range(n_nodes)]
            fed = EmulatedFederation(trainers, dsets, agg=agg)
            fed.fit(a.epochs)
            sc = score(trainers[0], ds_all, all_thetas, cfg["n_docs_global_inf"])  # every client holds the global model
            rows.append(dict(arm=f"federated_{agg}", seed=seed, seconds=time.perf_counter() - t0,
                             global_steps=fed.global_step, **sc))
            print(rows[-1], flush=True)
            del fed, trainers
            torch.cuda.empty_cache()
        # ---- the unmodified reference on the union (same box)
        if a.reference:
            import logging

            from baseline.ref_arm import _import_reference

            RefAVITM, _, RefBOW, _, _, _ = _import_reference()
            ep = a.reference_epochs or a.epochs
            dense = np.asarray(all_counts.todense(), dtype=np.float32)
            t0 = time.perf_counter()
            ref = RefAVITM(logger=logging.getLogger("reference"), input_size=V, n_components=K, model_type="prodLDA",
                           hidden_sizes=(100, 100), activation="softplus", dropout=0.2, learn_priors=True, batch_size=64,
                           lr=2e-3, momentum=0.99, solver="adam", num_epochs=ep, reduce_on_plateau=False,
                           topic_prior_mean=0.0, topic_prior_variance=None, num_samples=5, num_data_loader_workers=0,
                           verbose=False)
            ref.fit(RefBOW(dense, vocab))
            torch.cuda.synchronize()
            betas = np.asarray(ref.get_topic_word_distribution(), dtype=np.float64)
            rows.append(dict(arm="reference_centralized", seed=seed, seconds=time.perf_counter() - t0, epochs=ep,
                             tss=metrics.tss(betas, topics), npmi=metrics.npmi_coherence(betas, all_counts),
                             diversity=metrics.topic_diversity(betas)))
            print(rows[-1], flush=True)
    out = {"config": {"nodes": n_nodes, "docs_per_node": a.docs, "vocab": V, "topics": K, "eta": 0.01,
                      "frozen_topics": cfg["frozen_topics"], "hidden": [100, 100], "batch": 64, "epochs": a.epochs},
           "published": {"tss_centralized": [8.679, 0.042], "tss_non_collaborative": [7.571, 0.048],
                         "tss_random": [3.564, 0.098], "dss_centralized": [2555.5, 37.6],
                         "dss_non_collaborative": [3066.7, 14.0],
                         "source": "experiments/dss_tss/results/eta_variable/results.pickle (reference), Eta = 0.01"},
           "rows": rows}
    if a.out:
        os.makedirs(os.path.dirname(os.path.abspath(a.out)), exist_ok=True)
        with open(a.out, "w") as f:
            json.dump(out, f, indent=1)
    print(json.dumps(out["rows"], indent=1))


if __name__ == "__main__":
    main()
