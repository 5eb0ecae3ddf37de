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
    a = ap.parse_args(argv)
    files = sorted(glob.glob(os.path.join(a.wordlists, "*.json")))
    if a.use is not None:
        files = [f for f in files if os.path.splitext(os.path.basename(f))[0] in a.use]
    df = pd.read_parquet(a.source)
    out = TMWrapper().preproc_corpus_tm(df, text_col=a.text_col, stw_files=[f for f in files if "equiv" not in f],
                                        eq_files=[f for f in files if "equiv" in f], min_lemas=a.min_lemas,
                                        no_below=a.no_below, no_above=a.no_above, keep_n=a.keep_n, out_path=a.out)
    print(f"{len(df)} -> {len(out)} documents written to {a.out}")


if __name__ == "__main__":
    main()
