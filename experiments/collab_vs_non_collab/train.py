#!/usr/bin/env python
"""Collaborative vs non-collaborative comparison on a real corpus split by field of study.
Ref: experiments/collab_vs_non_collab/train.py:22-158 — trains one centralized model on the whole corpus
and one model per ``fos`` value, then reports NPMI coherence (against the full corpus), inverted RBO and
topic diversity for each."""
import argparse
import os
import sys

import numpy as np
import pandas as pd

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from gfedntm_b200.aux_modules.tm_wrapper import TMWrapper  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--source", required=True, help="preprocessed parquet with bow_text, fos[, embeddings]")
    ap.add_argument("--trainer", default="prod", choices=["prod", "lda", "ctm", "zeroshot"])
    ap.add_argument("--ntopics", type=int, default=50)
    ap.add_argument("--epochs", type=int, default=100)
    ap.add_argument("--out", default="collab_vs_non_collab.csv")
    a = ap.parse_args(argv)
    df = pd.read_parquet(a.source)
    tw = TMWrapper()

    def emb(frame):
        return np.stack(frame["embeddings"].values).astype(np.float32) if "embeddings" in frame else None

    rows = []
    central, ds, secs = tw.train_root_model(df["bow_text"].tolist(), a.trainer, a.ntopics, emb(df), num_epochs=a.epochs)
    ref_bow = ds.X
    rows.append({"model": "centralized", "n_docs": len(df), "seconds": secs,
                 "npmi": tw.calculate_cohr_vs_ref(central, ref_bow), "rbo": tw.calculate_rbo(central),
                 "td": tw.calculate_td(central)})
    for fos, part in df.groupby("fos"):
        if len(part) < 20:
            continue
        m, d, secs = tw.train_root_model(part["bow_text"].tolist(), a.trainer, a.ntopics, emb(part), num_epochs=a.epochs)
        rows.append({"model": f"non_collab:{fos}", "n_docs": len(part), "seconds": secs,
                     "npmi": tw.calculate_cohr_vs_ref(m, d.X), "rbo": tw.calculate_rbo(m), "td": tw.calculate_td(m)})
    out = pd.DataFrame(rows)
    out.to_csv(a.out, index=False)
    print(out)


if __name__ == "__main__":
    main()
