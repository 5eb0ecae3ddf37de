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
