#!/usr/bin/env python
"""Corpus preprocessing CLI.  Ref: aux_scripts/preprocessing/text_preproc.py:15-171 (interactive choice of
stop-word / equivalence lists, then the topicmodeler preprocessing pipeline).  This version takes the
lists on the command line and runs gfedntm_b200's in-process pipeline (no Dask / Spark)."""
import argparse
import glob
import os
import sys

import pandas as pd

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__)))))
from gfedntm_b200.aux_modules.tm_wrapper import TMWrapper  # noqa: E402


def main(argv=None):
    ap = argparse.ArgumentParser(description=__doc__)
    ap.add_argument("--source", required=True, help="parquet file/dir with the raw corpus")
    ap.add_argument("--out", required=True, help="output parquet")
    ap.add_argument("--text_col", default="lemmas")
    ap.add_argument("--wordlists", default=os.path.join(os.path.dirname(__file__), "wordlists"),
                    help="directory with *_stopwords.json / *_equivalences.json lists")
    ap.add_argument("--use", nargs="*", default=None, help="list names to apply (default: all in the directory)")
    ap.add_argument("--min_lemas", type=int, default=3)
    ap.add_argument("--no_below", type=int, default=1)
    ap.add_argument("--no_above", type=float, default=1.0)
    ap.add_argument("--keep_n", type=int, default=100000)
    ap.add_argument("--interactive", action="store_true",
                    help="choose the stop-word and equivalence lists at a prompt, like the reference's script")
    a = ap.parse_args(argv)
    files = sorted(glob.glob(os.path.join(a.wordlists, "*.json")))
    # label lists (valid_for = "labels", e.g. wiki_categories) are not preprocessing lists
    import json

    kinds = {f: json.load(open(f, encoding="utf-8")).get("valid_for", "stopwords") for f in files}
    files = [f for f in files if kinds[f] in ("stopwords", "equivalences")]
    if a.interactive:
        from gfedntm_b200.aux_modules.misc import var_num_keyboard

        chosen = []
        for kind in ("stopwords", "equivalences"):
            cand = [f for f in files if kinds[f] == kind]
            print(f"\nAvailable {kind} lists:")
            for i, f in enumerate(cand):
                print(f"  [{i}] {os.path.splitext(os.path.basename(f))[0]}")
            raw = input(f"Numbers of the {kind} lists to apply (comma separated, empty = none): ").strip()
            chosen += [cand[int(t)] for t in raw.split(",") if t.strip().isdigit() and int(t) < len(cand)]
        a.min_lemas = var_num_keyboard("int", a.min_lemas, "Minimum number of lemmas per document")
        a.no_below = var_num_keyboard("int", a.no_below, "Minimum number of documents a term must appear in")
        a.no_above = var_num_keyboard("float", a.no_above, "Maximum fraction of documents a term may appear in")
        a.keep_n = var_num_keyboard("int", a.keep_n, "Maximum vocabulary size")
        files = chosen
    elif a.use is not None:
        files = [f for f in files if os.path.splitext(os.path.basename(f))[0] in a.use]
    df = pd.read_parquet(a.source)
    out = TMWrapper().preproc_corpus_tm(df, text_col=a.text_col, stw_files=[f for f in files if kinds[f] == "stopwords"],
                                        eq_files=[f for f in files if kinds[f] == "equivalences"], min_lemas=a.min_lemas,
                                        no_below=a.no_below, no_above=a.no_above, keep_n=a.keep_n, out_path=a.out)
    print(f"{len(df)} -> {len(out)} documents written to {a.out}")


if __name__ == "__main__":
    main()
