#!/usr/bin/env python
"""Command-line front end for preprocessing / centralized training / second-level models, driven by a model
folder with a ``config.json`` -- the calling convention of the tool the reference's TMWrapper shells out to
(``topicmodeling.py --preproc|--train|--hierarchical --config <file> [--config_child <file>] [--nw N]``,
src/aux_modules/tmWrapper/tm_wrapper.py:123-169,345 -> topicmodeler/src/topicmodeling/topicmodeling.py:1836-2212).

Everything runs in process on gfedntm_b200's own pipeline and trainers; the Java / Spark trainers of the original
tool (``mallet``, ``sparkLDA``) are not available on the target boxes and are refused with a clear message.

``config.json`` (keys of the reference's format that are used):

    {"name": ..., "trainer": "prodLDA" | "ctm" | "zeroshot" | "neuralLDA",
     "TrDtSet": "<training set json>"            # {"Dtsets": [{"parquet": ..., "idfld": ..., "lemmasfld": [...]}]}
     "Preproc": {"min_lemas", "no_below", "no_above", "keep_n", "stopwords": [...], "equivalences": [...]},
     "TMparam": {"ntopics", "num_epochs", "batch_size", "hidden_sizes", "lr", ...},
     "hierarchy-level": 0 | 1, "htm-version": "htm-ws" | "htm-ds", "expansion_tpc": k, "thr": t}   # child models

Outputs in the model folder: ``corpus.parquet`` (--preproc), ``TMmodel/`` (``aux_modules/tm_model.py`` layout) and
``model.pth`` (--train / --hierarchical).
"""
from __future__ import annotations

import argparse
import json
import logging
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parents[2]
if str(ROOT) not in sys.path:
    sys.path.insert(0, str(ROOT))

LOG = logging.getLogger("topicmodeling")
UNAVAILABLE = {"mallet": "Mallet (Java) is not shipped", "sparklda": "Spark is not available on this box"}


def _wordlist_paths(names, folder: Path):
    """List names (with or without .json) -> files: next to the config, then the shipped word-list folders."""
    search = [folder, folder / "wordlists", ROOT / "aux_scripts" / "preprocessing" / "wordlists", ROOT / "static" / "wordlists"]
    out = []
    for n in names or []:
        n = n if n.endswith(".json") else n + ".json"
        for d in search:
            if (d / n).exists():
                out.append(str(d / n))
                break
        else:
            raise FileNotFoundError(f"word list {n} not found in {[str(d) for d in search]}")
    return out


def _load_config(path: str):
    cfg_path = Path(path).resolve()
    with open(cfg_path, encoding="utf-8") as f:
        cfg = json.load(f)
    return cfg, cfg_path.parent


def _training_frames(cfg: dict, folder: Path):
    import pandas as pd

    ds_file = cfg.get("TrDtSet")
    if ds_file is None:
        raise ValueError("config.json has no TrDtSet")
    p = Path(ds_file)
    p = p if p.is_absolute() else (folder / p if (folder / p).exists() else ROOT / p)
    with open(p, encoding="utf-8") as f:
        tr = json.load(f)
    frames = []
    for d in tr["Dtsets"]:
        pq = Path(d["parquet"])
        pq = pq if pq.is_absolute() else (folder / pq if (folder / pq).exists() else ROOT / pq)
        df = pd.read_parquet(pq)
        if d.get("filter"):
            df = df.query(d["filter"])
        lem = d.get("lemmasfld") or ["lemmas"]
        out = pd.DataFrame({"id": df[d.get("idfld", "id")].astype(str).values,
                            "lemmas": df[lem].astype(str).agg(" ".join, axis=1).values})
        for col in ("embeddings", "fieldsOfStudy", "fos"):
            if col in df:
                out[col] = df[col].values
        frames.append(out)
    return pd.concat(frames, ignore_index=True)


def do_preproc(cfg: dict, folder: Path, nw: int = 0):
    from gfedntm_b200.aux_modules.tm_wrapper import TMWrapper

    pre = cfg.get("Preproc", {})
    df = _training_frames(cfg, folder)
    out = TMWrapper(LOG).preproc_corpus_tm(
        df, text_col="lemmas", stw_files=_wordlist_paths(pre.get("stopwords"), folder),
        eq_files=_wordlist_paths(pre.get("equivalences"), folder), min_lemas=int(pre.get("min_lemas", 3)),
        no_below=int(pre.get("no_below", 1)), no_above=float(pre.get("no_above", 1.0)), keep_n=int(pre.get("keep_n", 100000)),
        out_path=str(folder / "corpus.parquet"))
    vocab = sorted({t for doc in out["bow_text"] for t in doc.split()})
    (folder / "vocabulary.txt").write_text("\n".join(vocab) + "\n", encoding="utf-8")
    LOG.info("preprocessed %d -> %d documents, %d terms", len(df), len(out), len(vocab))
    return out


def _embeddings(df):
    if "embeddings" not in df:
        return None
    e = df["embeddings"].values
    if len(e) and isinstance(e[0], str):
        return np.array([np.array(x.split(), dtype=np.float32) for x in e])
    return np.stack([np.asarray(x, dtype=np.float32) for x in e])


def do_train(cfg: dict, folder: Path, father=None):
    import pandas as pd
    import torch

    from gfedntm_b200.aux_modules.tm_model import TMmodel
    from gfedntm_b200.aux_modules.tm_wrapper import TMWrapper

    trainer = str(cfg.get("trainer", "prodLDA")).lower()
    if trainer in UNAVAILABLE:
        raise SystemExit(f"trainer '{trainer}' is unavailable here: {UNAVAILABLE[trainer]}; use prodLDA / neuralLDA / ctm / zeroshot")
    corpus = folder / "corpus.parquet"
    if not corpus.exists():
        do_preproc(cfg, folder)
    df = pd.read_parquet(corpus)
    tm = dict(cfg.get("TMparam", {}))
    ntopics = int(tm.pop("ntopics", tm.pop("n_components", 50)))
    epochs = int(tm.pop("num_epochs", 100))
    kw = {k: (tuple(v) if k == "hidden_sizes" else v) for k, v in tm.items()
          if k in ("batch_size", "hidden_sizes", "lr", "momentum", "dropout", "activation", "solver", "learn_priors", "seed")}
    name = {"prodlda": "prod", "prod": "prod", "neurallda": "lda", "lda": "lda", "ctm": "ctm", "combined": "ctm",
            "zeroshot": "zeroshot"}.get(trainer)
    if name is None:
        raise SystemExit(f"unknown trainer '{trainer}'")
    w = TMWrapper(LOG)
    model, ds, secs = w.train_root_model(df["bow_text"].tolist(), trainer=name, ntopics=ntopics, embeddings=_embeddings(df),
                                         max_features=int(tm.get("max_features", 100000)), num_epochs=epochs, **kw)
    betas = np.asarray(model.get_topic_word_distribution())
    thetas = np.asarray(model.get_doc_topic_distribution(ds, n_samples=int(tm.get("num_samples", 20))))
    vocab = list(ds.idx2token)
    bow = ds.X if hasattr(ds, "X") else ds.X_bow
    TMmodel(str(folder / "TMmodel")).create(betas=betas, thetas=thetas, vocab=vocab, reference_bow=bow)
    torch.save({"state_dict": model.model.state_dict(), "vocab": vocab, "trainer": trainer, "ntopics": ntopics},
               folder / "model.pth")
    (folder / "train_report.json").write_text(json.dumps(
        {"trainer": trainer, "ntopics": ntopics, "epochs": epochs, "documents": len(df), "vocabulary": len(vocab),
         "seconds": secs, "backend": model.backend}, indent=1))
    LOG.info("trained %s with %d topics on %d documents in %.1f s (%s backend)", trainer, ntopics, len(df), secs, model.backend)
    return model, ds


def do_hierarchical(cfg_father: dict, folder_father: Path, cfg_child: dict, folder_child: Path):
    """Second-level model: builds the child's corpus from the father's model (HTM-WS / HTM-DS) and trains it."""
    import torch

    from gfedntm_b200.aux_modules.tm_model import TMmodel
    from gfedntm_b200.aux_modules.tm_wrapper import TMWrapper

    trainer = str(cfg_child.get("trainer", cfg_father.get("trainer", "prodLDA"))).lower()
    if trainer in UNAVAILABLE:
        raise SystemExit(f"trainer '{trainer}' is unavailable here: {UNAVAILABLE[trainer]}")
    father, fds = do_train(cfg_father, folder_father) if not (folder_father / "model.pth").exists() else _reload(cfg_father, folder_father)
    tm = dict(cfg_child.get("TMparam", {}))
    name = {"prodlda": "prod", "prod": "prod", "neurallda": "lda", "lda": "lda", "ctm": "ctm", "zeroshot": "zeroshot"}[trainer]
    version = str(cfg_child.get("htm-version", "htm-ds")).upper()
    m, ds, keep, secs = TMWrapper(LOG).train_htm_submodel(
        version, father, fds, trainer=name, expansion_topic=int(cfg_child.get("expansion_tpc", 0)),
        thr=cfg_child.get("thr"), ntopics=int(tm.get("ntopics", 10)), num_epochs=int(tm.get("num_epochs", 100)))
    betas = np.asarray(m.get_topic_word_distribution())
    thetas = np.asarray(m.get_doc_topic_distribution(ds, n_samples=5))
    bow = ds.X if hasattr(ds, "X") else ds.X_bow
    folder_child.mkdir(parents=True, exist_ok=True)
    TMmodel(str(folder_child / "TMmodel")).create(betas=betas, thetas=thetas, vocab=list(ds.idx2token), reference_bow=bow)
    torch.save({"state_dict": m.model.state_dict(), "vocab": list(ds.idx2token), "kept_documents": np.asarray(keep)},
               folder_child / "model.pth")
    LOG.info("%s child of topic %s: %d documents, %.1f s", version, cfg_child.get("expansion_tpc", 0), len(keep), secs)


def _reload(cfg: dict, folder: Path):
    # (training is fast on this back end: re-train the father rather than re-implementing model restoration here)
    return do_train(cfg, folder)


def main(argv=None):
    ap = argparse.ArgumentParser(description="preprocessing / training CLI of gfedntm_b200 (config.json driven)")
    ap.add_argument("--preproc", action="store_true")
    ap.add_argument("--train", action="store_true")
    ap.add_argument("--hierarchical", action="store_true")
    ap.add_argument("--spark", action="store_true", help="accepted for compatibility; Spark is unavailable")
    ap.add_argument("--config", required=True)
    ap.add_argument("--config_child", default=None)
    ap.add_argument("--nw", type=int, default=0, help="workers (accepted for compatibility; the pipeline is in process)")
    a = ap.parse_args(argv)
    logging.basicConfig(level=logging.INFO, format="%(asctime)s [%(levelname)-5.5s] %(message)s")
    if a.spark:
        raise SystemExit("--spark: Spark is not available on this box; run without it")
    cfg, folder = _load_config(a.config)
    t0 = time.perf_counter()
    if a.preproc:
        do_preproc(cfg, folder, a.nw)
    if a.train:
        do_train(cfg, folder)
    if a.hierarchical:
        if not a.config_child:
            raise SystemExit("--hierarchical needs --config_child")
        cfg_c, folder_c = _load_config(a.config_child)
        do_hierarchical(cfg, folder, cfg_c, folder_c)
    if not (a.preproc or a.train or a.hierarchical):
        raise SystemExit("nothing to do: pass --preproc, --train or --hierarchical")
    LOG.info("done in %.1f s", time.perf_counter() - t0)


if __name__ == "__main__":
    main()
