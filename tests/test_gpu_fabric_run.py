"""`main.py --fabric N` (in-box federation, one client per GPU) and mid-training checkpoints of the fused engine.

Ref layout: static/output_models/client<id>/model_<id>_<date>.npz and server/global_model_<date>.npz with the keys
betas / thetas / ntopics / topics (auxiliary_functions.py:66-99, federated_model.py:151-197)."""
import configparser
import os
import subprocess
import sys

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fabric_checkpoint_roundtrip_on_the_fused_engine(tmp_path):
    """Save after 3 steps, take 2 more; a fresh trainer restored from the file must reproduce those 2 steps
    (weights, Adam moments, step counters, Philox counters and the loader's shuffle state all come back)."""
    from gfedntm_b200.federated.trainer import FederatedTrainer
    from gfedntm_b200.models import AVITM, BOWDataset
    from gfedntm_b200.utils.checkpoint import load_fabric_checkpoint, save_fabric_checkpoint

    rng = np.random.default_rng(0)
    X = rng.poisson(0.05, (320, 1200)).astype(np.float32)
    X[X.sum(1) == 0, 0] = 1

    def make():
        model = AVITM(input_size=1200, n_components=12, hidden_sizes=(32, 32), batch_size=64, num_epochs=1,
                      verbose=False, device="cuda:0", seed=4)
        tr = FederatedTrainer(model, n_samples=320, world_size=1, rank=0)
        loader = model.make_loader(BOWDataset(X, None), shuffle=True)
        tr.prepare(loader)
        return model, tr, loader

    def run(tr, loader, n, it=None):
        it = it if it is not None else iter(loader)
        out = []
        for _ in range(n):
            try:
                mb = next(it)
            except StopIteration:
                it = iter(loader)
                mb = next(it)
            out.append(float(tr.step(mb.bow, mb.x_ctx).item()))
        return out, it

    model, tr, loader = make()
    _, it = run(tr, loader, 3)
    if hasattr(it, "close"):
        it.close()
    save_fabric_checkpoint(str(tmp_path), tr, loader, global_step=3)
    want, it = run(tr, loader, 2)
    want_beta = model.model.beta.detach().clone()
    want_m = model.engine.flat.M.detach().clone()

    model2, tr2, loader2 = make()
    info = load_fabric_checkpoint(str(tmp_path), tr2, loader2)
    assert info["global_step"] == 3 and model2.engine.step_count == 3
    got, _ = run(tr2, loader2, 2)
    np.testing.assert_allclose(got, want, rtol=1e-4)
    torch.testing.assert_close(model2.model.beta.detach(), want_beta, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(model2.engine.flat.M, want_m, rtol=1e-2, atol=1e-4)  # (fp32 atomics reorder sums)


def _write_config(tmp_path, extra=None):
    cfg = configparser.ConfigParser()
    cfg.read(os.path.join(ROOT, "config", "dft_params.cf"))
    cfg.set("ntms", "n_components", "6")
    cfg.set("ntms", "hidden_sizes", "(16, 16)")
    cfg.set("ntms", "num_epochs", "2")
    cfg.set("ntms", "batch_size", "32")
    cfg.set("ntms", "num_samples", "3")
    cfg.set("b200", "agg", "grad")
    cfg.set("b200", "backend", "fused")
    for k in ("save_client", "save_server", "logs_client", "logs_server"):
        cfg.set("save_dir", k, str(tmp_path / "out" / k.replace("save_", "models_")))
    for sec, key, val in (extra or []):
        cfg.set(sec, key, val)
    path = str(tmp_path / "params.cf")
    with open(path, "w") as f:
        cfg.write(f)
    return path


@pytest.mark.parametrize("share", ["all", "subset"])
def test_main_fabric_two_clients_npz_layout_and_resume(tmp_path, share):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs at least 2 GPUs")
    from gfedntm_b200.utils.synthetic import generate_federated_corpus, save_synthetic_npz

    topics, doc_topics, counts = generate_federated_corpus(vocab_size=400, n_topics=6, n_docs=96, n_nodes=2,
                                                           frozen_topics=1, nwords=(30, 50), seed=2)
    npz = str(tmp_path / "synthetic.npz")
    save_synthetic_npz(npz, topics, doc_topics, counts)
    extra = [("federation", "grads_to_share", "prior_mean,prior_variance,beta")] if share == "subset" else None
    cfg = _write_config(tmp_path, extra)
    ckpt = str(tmp_path / "ckpt")
    env = dict(os.environ, PYTHONPATH=ROOT)
    base = [sys.executable, os.path.join(ROOT, "main.py"), "--config", cfg, "--source", npz, "--data_type", "synthetic",
            "--model_type", "avitm", "--fabric", "2", "--checkpoint_dir", ckpt]
    res = subprocess.run(base + ["--max_iters", "4", "--checkpoint_every", "2"], env=env, capture_output=True, text=True,
                         timeout=600)
    assert res.returncode == 0, res.stdout[-3000:] + res.stderr[-3000:]
    assert os.path.exists(os.path.join(ckpt, "fabric_rank0.ckpt")) and os.path.exists(os.path.join(ckpt, "fabric_rank1.ckpt"))
    out = tmp_path / "out"
    betas = []
    for cid in (1, 2):
        files = list(out.glob(f"models_client{cid}/model_{cid}_*.npz"))
        assert files, [str(p) for p in out.rglob("*")]
        z = np.load(str(files[0]), allow_pickle=True)
        assert set(z.files) >= {"betas", "thetas", "ntopics", "topics"}
        V = z["betas"].shape[1]
        assert z["betas"].shape[0] == 6 and z["thetas"].shape == (96, 6) and int(z["ntopics"]) == 6
        np.testing.assert_allclose(z["betas"].sum(1), np.ones(6), rtol=1e-4)
        np.testing.assert_allclose(z["thetas"].sum(1), np.ones(96), rtol=1e-6)
        assert np.asarray(z["topics"]).shape == (6, 10)
        betas.append(z["betas"])
    # beta is shared in both settings: both clients (and the server file) hold the same topic-word distribution
    np.testing.assert_allclose(betas[0], betas[1], rtol=1e-5, atol=1e-8)
    g = list((out / "models_server").glob("global_model_*.npz"))
    assert g
    zg = np.load(str(g[0]), allow_pickle=True)
    np.testing.assert_allclose(zg["betas"], betas[0], rtol=1e-5, atol=1e-8)
    # resume: continues from global step 2 of the checkpoint and finishes the remaining steps
    res2 = subprocess.run(base + ["--max_iters", "6", "--resume"], env=env, capture_output=True, text=True, timeout=600)
    assert res2.returncode == 0, res2.stdout[-3000:] + res2.stderr[-3000:]
    assert "resumed from" in (res2.stdout + res2.stderr)
