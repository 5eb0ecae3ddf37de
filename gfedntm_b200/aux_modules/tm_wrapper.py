"""In-process topic-model wrapper: preprocessing, centralized training and evaluation.

Ref: src/aux_modules/tmWrapper/tm_wrapper.py:1-400.  The reference shells out to the vendored
IntelComp ``topicmodeler`` (Dask/Spark preprocessing, Mallet, gensim coherence).  None of that is on the
north-star surface or installed on the target boxes; this wrapper offers the same entry points
(``preproc_corpus_tm``, ``train_root_model``, ``calculate_cohr_vs_ref``, ``calculate_rbo``,
``calculate_td``) implemented with gfedntm_b200's own trainers and metrics."""
from __future__ import annotations

import json
import logging
import time
from pathlib import Path
from typing import Optional, Sequence

import numpy as np
import pandas as pd
from sklearn.feature_extraction.text import CountVectorizer

from ..models import AVITM, CTM, BOWDataset, CTMDataset
from ..utils import metrics


def load_wordlists(paths: Sequence[str]):
    """Stop-word / equivalence lists in the reference's JSON format
    (``{"name", "description", "valid_for", "visibility", "wordlist": [...]}``; equivalences as ``a:b``)."""
    stop, equiv = set(), {}
    for p in paths:
        with open(p, encoding="utf-8") as f:
            d = json.load(f)
        for w in d.get("wordlist", []):
            if d.get("valid_for") == "equivalences" or ":" in w:
                a, _, b = w.partition(":")
                equiv[a.strip()] = b.strip()
            else:
                stop.add(w.strip())
    return stop, equiv


class TMWrapper:
    def __init__(self, logger: Optional[logging.Logger] = None) -> None:
        self._logger = logger or logging.getLogger("TMWrapper")

    # -- preprocessing ---------------------------------------------------------------------------
    def preproc_corpus_tm(self, df: pd.DataFrame, text_col: str = "lemmas", stw_files: Sequence[str] = (),
                          eq_files: Sequence[str] = (), min_lemas: int = 3, no_below: int = 1,
                          no_above: float = 1.0, keep_n: int = 100000, out_path: Optional[str] = None
                          ) -> pd.DataFrame:
        """Stop-word removal, equivalence replacement, minimum-length filter and vocabulary pruning
        (``no_below`` / ``no_above`` / ``keep_n`` as in gensim's ``filter_extremes``).  Returns a frame with
        ``id``, ``bow_text`` (+ ``embeddings`` / ``fos`` when present) and optionally writes parquet."""
        stop, equiv = load_wordlists(list(stw_files) + list(eq_files))
        docs = []
        for text in df[text_col].astype(str):
            toks = [equiv.get(t, t) for t in text.split()]
            docs.append([t for t in toks if t not in stop])
        cv = CountVectorizer(tokenizer=lambda s: s.split(), lowercase=False, token_pattern=None,
                             min_df=no_below, max_df=no_above, max_features=keep_n)
        cv.fit([" ".join(d) for d in docs])
        vocab = set(cv.vocabulary_)
        docs = [[t for t in d if t in vocab] for d in docs]
        keep = [len(d) >= min_lemas for d in docs]
        out = pd.DataFrame({"id": df["id"].values if "id" in df else np.arange(len(df)),
                            "bow_text": [" ".join(d) for d in docs]})
        for col in ("embeddings", "fos", "fieldsOfStudy"):
            if col in df:
                out[col if col != "fieldsOfStudy" else "fos"] = df[col].values
        out = out[np.asarray(keep)].reset_index(drop=True)
        if out_path:
            Path(out_path).parent.mkdir(parents=True, exist_ok=True)
            out.to_parquet(out_path)
        return out

    # -- training --------------------------------------------------------------------------------
    def train_root_model(self, corpus: Sequence[str], trainer: str = "prod", ntopics: int = 50,
                         embeddings: Optional[np.ndarray] = None, max_features: Optional[int] = 2000,
                         num_epochs: int = 100, **kw):
        """Centralized training of ProdLDA / NeuralLDA ('prod' / 'lda') or CTM ('ctm' / 'zeroshot').
        Returns ``(model, dataset, seconds)``."""
        t0 = time.perf_counter()
        cv = CountVectorizer(lowercase=True, stop_words="english", max_features=max_features)
        bow = cv.fit_transform(corpus)
        vocab = cv.get_feature_names_out()
        if trainer in ("prod", "prodlda", "lda"):
            ds = BOWDataset(bow, vocab)
            m = AVITM(input_size=len(vocab), n_components=ntopics,
                      model_type="LDA" if trainer == "lda" else "prodLDA", num_epochs=num_epochs,
                      verbose=False, **kw)
        else:
            if embeddings is None:
                raise ValueError("CTM training needs document embeddings")
            ds = CTMDataset(embeddings, bow, vocab)
            m = CTM(input_size=len(vocab), contextual_size=int(embeddings.shape[1]),
                    inference_type="zeroshot" if trainer == "zeroshot" else "combined", n_components=ntopics,
                    num_epochs=num_epochs, verbose=False, **kw)
        m.fit(ds)
        return m, ds, time.perf_counter() - t0

    def train_htm_submodel(self, version: str, father_model, father_dataset, trainer: str = "prod",
                           expansion_topic: int = 0, thr: Optional[float] = None, ntopics: int = 10,
                           num_epochs: int = 100, n_samples: int = 20, sample: bool = False,
                           seed: Optional[int] = None, **kw):
        """Second-level (hierarchical) model for ``expansion_topic`` of ``father_model`` (ref:
        tm_wrapper.py:279-355 -> topicmodeling.py:1562-1830).  ``version`` is ``"HTM-WS"`` (keep the word
        occurrences assigned to the topic) or ``"HTM-DS"`` (keep the documents whose topic weight is at least
        ``thr``).  The child is trained on the sub-corpus over the father's vocabulary.
        Returns ``(model, dataset, kept_document_indices, seconds)``."""
        from .hierarchical import htm_ds, htm_ws

        t0 = time.perf_counter()
        thetas = np.asarray(father_model.get_doc_topic_distribution(father_dataset, n_samples=n_samples))
        bow = father_dataset.X if hasattr(father_dataset, "X") else father_dataset.X_bow
        if version.upper() == "HTM-DS":
            sub, keep = htm_ds(bow, thetas, expansion_topic, 0.1 if thr is None else thr)
        elif version.upper() == "HTM-WS":
            betas = np.asarray(father_model.get_topic_word_distribution())
            sub, keep = htm_ws(bow, thetas, betas, expansion_topic, sample=sample, seed=seed)
        else:
            raise ValueError("version must be 'HTM-WS' or 'HTM-DS'")
        if sub.shape[0] < 2:
            raise ValueError(f"the expansion of topic {expansion_topic} keeps {sub.shape[0]} documents")
        vocab = father_dataset.idx2token
        self._logger.info("%s expansion of topic %d: %d of %d documents, %d word occurrences", version,
                          expansion_topic, sub.shape[0], bow.shape[0], int(sub.sum()))
        if trainer in ("prod", "prodlda", "lda"):
            ds = BOWDataset(sub.astype(np.float32), vocab)
            m = AVITM(input_size=sub.shape[1], n_components=ntopics,
                      model_type="LDA" if trainer == "lda" else "prodLDA", num_epochs=num_epochs, verbose=False,
                      **kw)
        else:
            emb = np.asarray(father_dataset.X_contextual)[keep]
            ds = CTMDataset(emb, sub.astype(np.float32), vocab)
            m = CTM(input_size=sub.shape[1], contextual_size=int(emb.shape[1]),
                    inference_type="zeroshot" if trainer == "zeroshot" else "combined", n_components=ntopics,
                    num_epochs=num_epochs, verbose=False, **kw)
        m.fit(ds)
        return m, ds, keep, time.perf_counter() - t0

    # -- the reference's calling convention: a model folder with config.json, a CLI run as a subprocess ----------
    def create_model_folder(self, path: str, name: str, trainer: str, training_set: str, tm_params: Optional[dict] = None,
                            preproc: Optional[dict] = None, description: str = "", **extra) -> str:
        """Writes ``<path>/<name>/config.json`` in the layout ``topicmodeling.py`` reads (ref:
        tm_wrapper.py:60-121: ``_get_model_config`` / ``_do_preproc`` build the same file).  ``training_set``: json
        describing the parquet sources (``{"Dtsets": [{"parquet", "idfld", "lemmasfld"}]}``)."""
        folder = Path(path) / name
        folder.mkdir(parents=True, exist_ok=True)
        cfg = {"name": name, "description": description, "trainer": trainer, "TrDtSet": str(training_set),
               "Preproc": {**dict(min_lemas=3, no_below=1, no_above=1.0, keep_n=100000, stopwords=[], equivalences=[]),
                           **(preproc or {})},
               "TMparam": dict(tm_params or {}), "hierarchy-level": 0}
        cfg.update(extra)
        with open(folder / "config.json", "w", encoding="utf-8") as f:
            json.dump(cfg, f, indent=1)
        return str(folder / "config.json")

    def _run_cli(self, *args: str) -> float:
        """Runs the CLI in a subprocess (like the reference, tm_wrapper.py:123-169) and returns its wall time."""
        import subprocess
        import sys

        cmd = [sys.executable, "-m", "gfedntm_b200.aux_modules.topicmodeling", *args]
        self._logger.info("running %s", " ".join(cmd))
        t0 = time.perf_counter()
        res = subprocess.run(cmd, capture_output=True, text=True, cwd=str(Path(__file__).resolve().parents[2]))
        dt = time.perf_counter() - t0
        if res.returncode != 0:
            raise RuntimeError(f"topicmodeling CLI failed ({res.returncode}):\n{res.stdout[-2000:]}\n{res.stderr[-2000:]}")
        return dt

    def preproc_corpus_cli(self, config_file: str, nw: int = 0) -> float:
        return self._run_cli("--preproc", "--config", config_file, "--nw", str(nw))

    def train_model_cli(self, config_file: str) -> float:
        return self._run_cli("--train", "--config", config_file)

    def train_htm_submodel_cli(self, config_father: str, config_child: str) -> float:
        return self._run_cli("--hierarchical", "--config", config_father, "--config_child", config_child)

    # -- evaluation ------------------------------------------------------------------------------
    @staticmethod
    def calculate_cohr_vs_ref(model, reference_bow, k: int = 10) -> float:
        """NPMI coherence of the model's topics with co-occurrences counted on a reference corpus
        (document-term matrix over the same vocabulary)."""
        return metrics.npmi_coherence(model.get_topic_word_distribution(), reference_bow, k)

    @staticmethod
    def calculate_rbo(model, k: int = 10) -> float:
        return metrics.inverted_rbo(model.get_topic_word_distribution(), k)

    @staticmethod
    def calculate_td(model, k: int = 25) -> float:
        return metrics.topic_diversity(model.get_topic_word_distribution(), k)
