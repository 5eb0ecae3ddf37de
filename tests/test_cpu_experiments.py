"""Smoke tests of the side tools: synthetic generator CLI and the DSS/TSS simulation."""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_generate_synthetic_cli(tmp_path):
    from gfedntm_b200.utils.generate_synthetic import main

    path = main(["--vocab_size", "200", "--n_topics", "6", "--n_docs", "20", "--n_nodes", "2", "--frozen_topics", "2",
                 "--nwords", "10", "20", "--out_dir", str(tmp_path)])
    z = np.load(path, allow_pickle=True)
    assert len(z["documents"]) == 2 and len(z["documents"][0]) == 20 and z["topic_vectors"].shape == (6, 200)
    assert os.path.exists(str(tmp_path / "node1.txt")) and os.path.exists(str(tmp_path / "node2.npz"))
    first = open(str(tmp_path / "node1.txt")).readline().split()
    assert first[0] == "0" and first[1] == "0" and first[2].startswith("wd")


def test_dss_tss_simulation_tiny(tmp_path):
    sys.path.insert(0, os.path.join(ROOT, "experiments", "dss_tss"))
    import run_simulation as rs

    cfg = {"vocab_size": 150, "n_topics": 5, "beta": 0.05, "alpha": 0.2, "n_docs": 60, "n_docs_inf": 30,
           "n_docs_global_inf": 40, "nwords": [20, 30], "n_nodes": 2, "frozen_topics": 1, "own_topics": 2,
           "experiment": "frozen", "values": [1, 2], "iters": 1}
    df = rs.run(cfg, str(tmp_path / "res.pickle"), epochs=2, iters=1, device="cpu")
    assert len(df) == 2 and {"TSS_centralized", "DSS_non_colab", "TSS_baseline"} <= set(df.columns)
    assert (df["TSS_centralized"] > 0).all() and (df["TSS_centralized"] <= 5 + 1e-6).all()


def test_quality_comparison_script_tiny(tmp_path):
    """The ours-vs-reference quality script runs end to end on CPU; the reference arm is optional (its sources
    live in the git-ignored baseline/_ref) and must degrade to an 'unavailable' row, not to a crash."""
    import subprocess

    out = tmp_path / "q.json"
    cmd = [sys.executable, os.path.join(ROOT, "experiments", "quality", "compare_with_reference.py"), "--docs", "200",
           "--vocab", "120", "--topics", "4", "--epochs", "1", "--seeds", "1", "--device", "cpu", "--out", str(out)]
    res = subprocess.run(cmd, capture_output=True, text=True, timeout=600)
    assert res.returncode == 0, res.stdout[-2000:] + res.stderr[-2000:]
    rows = json.load(open(out))["rows"]
    arms = [r["arm"] for r in rows]
    assert arms[0] == "random topics" and arms[1] == "generating topics" and any(a.startswith("ours") for a in arms)
    gen = rows[1]
    assert abs(gen["tss"] - 4.0) < 1e-6  # the generating topics score the maximum against themselves
    ours = [r for r in rows if r["arm"].startswith("ours")][0]
    assert 0.0 < ours["tss"] <= 4.0 + 1e-6


def test_collab_vs_non_collab_tiny(tmp_path):
    import pandas as pd

    sys.path.insert(0, os.path.join(ROOT, "experiments", "collab_vs_non_collab"))
    import train as cvn

    rng = np.random.default_rng(0)
    vocab = [f"wd{i}" for i in range(60)]
    rows = []
    for i in range(90):
        fos = ["physics", "biology", "history"][i % 3]
        lo = (i % 3) * 20
        words = rng.choice(vocab[lo:lo + 20], size=25)
        rows.append({"id": i, "bow_text": " ".join(words), "fos": fos})
    src = tmp_path / "corpus.parquet"
    pd.DataFrame(rows).to_parquet(src)
    out = tmp_path / "res.csv"
    cvn.main(["--source", str(src), "--trainer", "prod", "--ntopics", "3", "--epochs", "1", "--out", str(out)])
    df = pd.read_csv(out)
    assert len(df) >= 2  # the centralized model plus at least one per-field model


def test_scalability_grid_tiny(tmp_path):
    import csv

    sys.path.insert(0, os.path.join(ROOT, "aux_scripts", "scalability"))
    import scalability

    out = scalability.main(["--docs", "60", "--topics", "2", "3", "--vocab", "80", "--epochs", "1", "--device", "cpu",
                            "--out", str(tmp_path / "s.csv")])
    rows = list(csv.DictReader(open(out)))
    assert len(rows) == 2 and {r["n_topics"] for r in rows} == {"2", "3"}
    assert all(float(r["train_s"]) > 0 and int(r["peak_rss"]) > 0 for r in rows)


def test_exact_word_movers_distance():
    """aux_scripts/evaluation/wmd.py: the exact transport distance is symmetric, zero on identical lists, never below
    the relaxed nearest-neighbour bound, and equals the closed form when both lists have one word."""
    import importlib.util
    import os

    import numpy as np

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    spec = importlib.util.spec_from_file_location("wmd_script", os.path.join(root, "aux_scripts", "evaluation", "wmd.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    rng = np.random.default_rng(0)
    vec = {f"w{i}": rng.standard_normal(6) for i in range(12)}
    a, b = [f"w{i}" for i in range(5)], [f"w{i}" for i in range(4, 11)]
    assert mod.wmd(a, a, vec) < 1e-9
    assert abs(mod.wmd(a, b, vec) - mod.wmd(b, a, vec)) < 1e-9
    assert mod.wmd(a, b, vec) >= mod.relaxed_wmd(a, b, vec) - 1e-9
    assert abs(mod.wmd(["w1"], ["w2"], vec) - np.linalg.norm(vec["w1"] - vec["w2"])) < 1e-9
    # moving half of the mass: {w1, w2} -> {w1, w3} costs 0.5 * |w2 - w3| at most
    assert mod.wmd(["w1", "w2"], ["w1", "w3"], vec) <= 0.5 * np.linalg.norm(vec["w2"] - vec["w3"]) + 1e-9
    assert np.isnan(mod.wmd(["zz"], a, vec))


def test_tmwrapper_cli_preproc_train_and_child_model(tmp_path):
    """The reference's calling convention (tm_wrapper.py:123-169,345): a model folder with config.json, the
    preprocessing / training / hierarchical CLI run as a SUBPROCESS, on the shipped sample parquet."""
    import json
    import os

    import numpy as np

    from gfedntm_b200.aux_modules.tm_wrapper import TMWrapper

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    trset = tmp_path / "trset.json"
    trset.write_text(json.dumps({"name": "t", "Dtsets": [{"parquet": os.path.join(root, "static/datasets/s2cs_tiny.parquet"),
                                                          "source": "s2cs_tiny", "idfld": "id", "lemmasfld": ["lemmas"],
                                                          "filter": ""}]}))
    w = TMWrapper()
    cfg = w.create_model_folder(str(tmp_path), "root", "prodLDA", str(trset),
                                tm_params=dict(ntopics=5, num_epochs=3, batch_size=64, hidden_sizes=[16, 16], num_samples=2),
                                preproc=dict(min_lemas=10, stopwords=["english_generic", "S2_stopwords"],
                                             equivalences=["S2_equivalences"]))
    t_pre = w.preproc_corpus_cli(cfg)
    assert t_pre > 0 and (tmp_path / "root" / "corpus.parquet").exists() and (tmp_path / "root" / "vocabulary.txt").exists()
    vocab = (tmp_path / "root" / "vocabulary.txt").read_text().split()
    assert "the" not in vocab and "tumour" not in vocab and len(vocab) > 50
    t_tr = w.train_model_cli(cfg)
    assert t_tr > 0 and (tmp_path / "root" / "TMmodel").is_dir() and (tmp_path / "root" / "model.pth").exists()
    rep = json.loads((tmp_path / "root" / "train_report.json").read_text())
    assert rep["ntopics"] == 5 and rep["documents"] > 300
    child = w.create_model_folder(str(tmp_path), "child", "prodLDA", str(trset),
                                  tm_params=dict(ntopics=3, num_epochs=2), **{"hierarchy-level": 1, "htm-version": "htm-ds",
                                                                              "expansion_tpc": 0, "thr": 0.05})
    w.train_htm_submodel_cli(cfg, child)
    assert (tmp_path / "child" / "TMmodel").is_dir()
    # unavailable third-party trainers are refused with a clear message
    bad = w.create_model_folder(str(tmp_path), "bad", "mallet", str(trset))
    import pytest

    with pytest.raises(RuntimeError, match="unavailable"):
        w.train_model_cli(bad)
