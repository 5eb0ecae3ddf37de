#!/usr/bin/env python
"""Regenerates the sample data shipped under ``static/`` and the word lists of the preprocessing CLI.

The reference ships a 334-document excerpt of a scholarly corpus (``static/datasets/s2cs_tiny.parquet``, columns
``id, lemmas, embeddings (192-d), year, fieldsOfStudy``), its preprocessed form (``preprocessing_example/
s2cs_tiny_preproc.parquet``, columns ``id, bow_text, fos, embeddings``, five ``fos`` values, plus vocabulary /
dictionary / config files), stop-word / equivalence lists (``{"name","description","valid_for","visibility",
"wordlist"}``, equivalences as ``"a:b"``) and a synthetic LDA corpus (``static/datasets/synthetic.npz``).  None of
those files is copied here: this script *synthesises* stand-ins with the same schema from small hand-written term
pools (documents are drawn from a five-field mixture model, embeddings are random projections of the topic
mixture), so that ``main.py --data_type real``, the preprocessing CLI and the examples have something to run on.

    python aux_scripts/make_sample_data.py [--out ROOT]
"""
from __future__ import annotations

import argparse
import json
import os

import numpy as np

FIELDS = {
    "computer_science": "algorithm network learning model data system software graph query compiler cache kernel thread "
                        "protocol cluster latency bandwidth gradient tensor encoder decoder token vector index schema "
                        "scheduler memory parallel distributed federated privacy encryption database transaction".split(),
    "medicine": "patient clinical trial treatment therapy disease cancer tumor dose cohort diagnosis surgery infection "
                "vaccine immune antibody blood cardiac symptom chronic outcome mortality placebo screening biopsy "
                "prognosis hospital nurse pathology receptor inflammation".split(),
    "physics": "quantum particle field energy photon electron spin lattice phase transition magnetic plasma laser "
               "scattering relativity gravity boson fermion entropy thermodynamic superconductor wave spectrum collider "
               "neutrino cosmology vacuum symmetry momentum oscillator".split(),
    "economics": "market price inflation trade labor wage policy tax growth capital investment bank credit debt fiscal "
                 "monetary demand supply consumer firm auction equilibrium incentive welfare subsidy tariff export "
                 "recession productivity portfolio risk".split(),
    "biology": "gene protein cell genome species evolution mutation enzyme bacteria virus rna dna sequence expression "
               "membrane neuron tissue organism ecology habitat population metabolism pathway receptor chromosome "
               "phylogeny fungus plant photosynthesis".split(),
}
COMMON = "study result method analysis approach paper show propose present use base new effect high low time rate level " \
         "different significant increase performance process structure function change important".split()
STOP = "the of and to in a is that for on with as by this we are be an from at which it can has have our these was were " \
       "or not but also more than their its between into using used such may both each other over under however".split()


def _docs(rng, n, lo=40, hi=90):
    names = list(FIELDS)
    pools = [FIELDS[f] for f in names]
    out = []
    proj = rng.standard_normal((len(names) + 1, 192)) / 3.0
    for i in range(n):
        f = int(rng.integers(len(names)))
        mix = rng.dirichlet([0.3] * len(names))
        mix[f] += 1.5
        mix /= mix.sum()
        L = int(rng.integers(lo, hi))
        words = []
        for _ in range(L):
            u = rng.random()
            if u < 0.18:
                words.append(STOP[int(rng.integers(len(STOP)))])
            elif u < 0.38:
                words.append(COMMON[int(rng.integers(len(COMMON)))])
            else:
                k = int(rng.choice(len(names), p=mix))
                pool = pools[k]
                # Zipf-like draw inside a field
                j = min(int(rng.zipf(1.6)) - 1, len(pool) - 1)
                words.append(pool[j])
        emb = np.concatenate([mix, [1.0]]) @ proj + 0.05 * rng.standard_normal(192)
        out.append(dict(id=f"doc{i:05d}", words=words, field=names[f], emb=emb.astype(np.float64),
                        year=int(rng.integers(1995, 2023))))
    return out


def _wordlist(name, description, valid_for, words):
    return {"name": name, "description": description, "valid_for": valid_for, "visibility": "Public", "wordlist": list(words)}


def make_wordlists(root):
    lists = {
        "english_generic.json": _wordlist("english_generic", "Frequent English function words", "stopwords",
                                          sorted(set(STOP + "about above after again all am any because been before being "
                                                            "below did do does doing down during few further had he her "
                                                            "here hers him his how i if just me most my no nor now off "
                                                            "once only out own same she should so some then there they "
                                                            "those through too until up very what when where while who "
                                                            "whom why will you your yours".split()))),
        "S2_stopwords.json": _wordlist("S2_stopwords", "Boilerplate terms of scholarly abstracts", "stopwords",
                                       sorted(set(COMMON + "abstract introduction conclusion et al fig figure table section "
                                                           "work article author copyright rights reserved elsevier springer "
                                                           "doi journal volume issue page keywords background objective "
                                                           "aim purpose finding".split()))),
        "S2CS_stopwords.json": _wordlist("S2CS_stopwords", "Boilerplate terms of computer-science abstracts", "stopwords",
                                         sorted(set("paper propose proposed novel approach framework experiment experimental "
                                                    "evaluation evaluate baseline benchmark state art outperform achieve "
                                                    "demonstrate implementation implement prototype scenario technique "
                                                    "problem solution task dataset result show".split()))),
        "AI_stopwords.json": _wordlist("AI", "Generic terms of artificial-intelligence texts", "stopwords",
                                       "ai intelligent intelligence artificial smart automated automatic application "
                                       "applications tool tools technology technologies solution solutions based "
                                       "driven enabled powered".split()),
        "cancer_stopwords.json": _wordlist("cancer_stopwords", "Generic terms of oncology project descriptions", "stopwords",
                                           "cancer patient patients study studies clinical research project aim aims "
                                           "novel develop development disease treatment health".split()),
        "cordis_stopwords.json": _wordlist("cordis_stopwords", "Generic terms of research-project summaries", "stopwords",
                                           "project projects european europe eu objective objectives work package "
                                           "partner partners consortium innovation innovative funding programme call "
                                           "deliverable impact".split()),
        "federated_stop.json": _wordlist("S2CS_stopwords for federated", "Stop words of the federated sample corpus",
                                         "stopwords", sorted(set(STOP + COMMON))),
        "S2_equivalences.json": _wordlist("S2_equivalences", "Spelling / plural normalisation for scholarly text",
                                          "equivalences",
                                          ["tumour:tumor", "modelling:modeling", "behaviour:behavior", "colour:color",
                                           "networks:network", "models:model", "algorithms:algorithm", "genes:gene",
                                           "proteins:protein", "cells:cell", "markets:market", "prices:price",
                                           "patients:patient", "particles:particle", "datasets:dataset", "data_set:dataset"]),
        "S2CS_equivalences.json": _wordlist("S2CS_equivalences", "Normalisation of computer-science terms", "equivalences",
                                            ["neural_networks:neural_network", "nn:neural_network", "gpus:gpu",
                                             "cpus:cpu", "databases:database", "db:database", "compilers:compiler",
                                             "ml:machine_learning", "dl:deep_learning", "nlp:natural_language_processing",
                                             "os:operating_system", "iot:internet_of_things"]),
        "AI_equivalences.json": _wordlist("AI_eqs", "Normalisation of artificial-intelligence terms", "equivalences",
                                          ["a.i.:ai", "artificial_intelligence:ai", "machine-learning:machine_learning",
                                           "deep-learning:deep_learning", "convnet:cnn", "convolutional_network:cnn",
                                           "recurrent_network:rnn", "transformers:transformer", "llms:llm"]),
        "cancer_equivalences.json": _wordlist("cancer_equivalences", "Normalisation of oncology terms", "equivalences",
                                              ["tumours:tumor", "tumors:tumor", "carcinomas:carcinoma",
                                               "metastases:metastasis", "chemo:chemotherapy", "radio:radiotherapy",
                                               "oncological:oncology", "biomarkers:biomarker"]),
        "cordis_equivalences.json": _wordlist("cordis_equivalences", "Normalisation of project-summary terms",
                                              "equivalences", ["smes:sme", "r&d:research_development", "e.u.:eu",
                                                               "co-ordination:coordination", "programmes:programme"]),
        "federated_equiv.json": _wordlist("S2_equivalences for federated corpus", "Equivalences of the federated sample corpus",
                                          "equivalences", ["networks:network", "models:model", "genes:gene", "cells:cell",
                                                           "markets:market", "patients:patient", "fields:field"]),
        "wiki_categories.json": _wordlist("wiki_categories", "Candidate topic labels", "labels",
                                          sorted({w.replace("_", " ").title() for f in FIELDS for w in FIELDS[f]} |
                                                 {f.replace("_", " ").title() for f in FIELDS})),
    }
    d = os.path.join(root, "aux_scripts", "preprocessing", "wordlists")
    os.makedirs(d, exist_ok=True)
    for fn, obj in lists.items():
        with open(os.path.join(d, fn), "w") as f:
            json.dump(obj, f, indent=1)
    ds = os.path.join(root, "static", "wordlists")
    os.makedirs(ds, exist_ok=True)
    for fn in ("S2_stopwords.json", "S2_equivalences.json"):
        with open(os.path.join(ds, fn), "w") as f:
            json.dump(lists[fn], f, indent=1)
    return lists


def main():
    import pandas as pd

    ap = argparse.ArgumentParser()
    ap.add_argument("--out", default=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    args = ap.parse_args()
    root = args.out
    rng = np.random.default_rng(20240607)
    lists = make_wordlists(root)
    docs = _docs(rng, 334)
    dset = os.path.join(root, "static", "datasets")
    os.makedirs(os.path.join(dset, "preprocessing_example"), exist_ok=True)
    raw = pd.DataFrame({"id": [d["id"] for d in docs], "lemmas": [" ".join(d["words"]) for d in docs],
                        "embeddings": [d["emb"] for d in docs], "year": [d["year"] for d in docs],
                        "fieldsOfStudy": [[d["field"]] for d in docs]})
    raw.to_parquet(os.path.join(dset, "s2cs_tiny.parquet"))
    # preprocessed form: stop words removed, equivalences applied, documents with fewer than 15 lemmas dropped
    stop = set(lists["english_generic.json"]["wordlist"]) | set(lists["S2_stopwords.json"]["wordlist"])
    eq = dict(e.split(":") for e in lists["S2_equivalences.json"]["wordlist"])
    rows = []
    for d in docs:
        toks = [eq.get(w, w) for w in d["words"] if w not in stop]
        if len(toks) >= 15:
            rows.append(dict(id=d["id"], bow_text=" ".join(toks), fos=d["field"],
                             embeddings=" ".join(f"{x:.6f}" for x in d["emb"])))
    rows = rows[:326]
    pre = pd.DataFrame(rows)
    pre.to_parquet(os.path.join(dset, "preprocessing_example", "s2cs_tiny_preproc.parquet"))
    vocab = {}
    for r in rows:
        for w in set(r["bow_text"].split()):
            vocab[w] = vocab.get(w, 0) + 1
    words = sorted(vocab)
    with open(os.path.join(dset, "preprocessing_example", "vocabulary.txt"), "w") as f:
        f.write("\n".join(words) + "\n")
    with open(os.path.join(dset, "preprocessing_example", "dictionary.gensim"), "w") as f:
        f.write(f"{len(rows)}\n")
        for i, w in enumerate(words):
            f.write(f"{i}\t{w}\t{vocab[w]}\n")
    cfg = {"name": "s2cs_tiny_preproc", "description": "synthetic stand-in of the preprocessing example",
           "Preproc": {"min_lemas": 15, "no_below": 1, "no_above": 1.0, "keep_n": 100000,
                       "stopwords": ["english_generic.json", "S2_stopwords.json"], "equivalences": ["S2_equivalences.json"]},
           "TrDtSet": "scholar_dtst.json", "trainer": "ctm"}
    with open(os.path.join(dset, "preprocessing_example", "config.json"), "w") as f:
        json.dump(cfg, f, indent=1)
    with open(os.path.join(dset, "preprocessing_example", "scholar_dtst.json"), "w") as f:
        json.dump({"name": "scholar_dtst", "Dtsets": [{"parquet": "static/datasets/s2cs_tiny.parquet", "source": "s2cs_tiny",
                                                       "idfld": "id", "lemmasfld": ["lemmas"], "filter": ""}]}, f, indent=1)
    # small synthetic LDA corpus in the layout main.py --data_type synthetic expects (5 nodes)
    from gfedntm_b200.utils.synthetic import generate_federated_corpus, save_synthetic_npz

    topics, doc_topics, counts = generate_federated_corpus(vocab_size=1000, n_topics=20, n_docs=200, n_nodes=5,
                                                           frozen_topics=3, nwords=(60, 100), seed=7)
    save_synthetic_npz(os.path.join(dset, "synthetic.npz"), topics, doc_topics, counts)
    print(f"wrote {len(docs)} raw / {len(rows)} preprocessed documents, {len(words)} vocabulary terms, "
          f"{len(lists)} word lists and synthetic.npz under {root}")


if __name__ == "__main__":
    import sys

    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    main()
