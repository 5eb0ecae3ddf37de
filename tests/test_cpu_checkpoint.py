"""Mid-training checkpoint / resume and the observability helpers."""
import json

import numpy as np
import torch


def _toy(n=96, v=120, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.poisson(0.1, (n, v)).astype(np.float32)
    X[X.sum(1) == 0, 0] = 1
    return X, [f"w{i}" for i in range(v)]


def test_checkpoint_resume_reproduces_training(tmp_path):
    from gfedntm_b200.federated import FederatedAVITM
    from gfedntm_b200.models import BOWDataset
    from gfedntm_b200.utils.checkpoint import load_training_checkpoint, save_training_checkpoint

    X, vocab = _toy()
    params = dict(input_size=120, n_components=4, model_type="prodLDA", hidden_sizes=(8, 8), activation="softplus",
                  dropout=0.0, learn_priors=True, batch_size=32, lr=2e-3, momentum=0.99, solver="adam",
                  num_epochs=5, reduce_on_plateau=False, topic_prior_mean=0.0, topic_prior_variance=None,
                  num_samples=2, num_data_loader_workers=0, verbose=False, seed=5)

    def make():
        torch.manual_seed(0)
        m = FederatedAVITM(params, ["beta"], None, backend="eager", device="cpu")
        m.model.inf_net.dropout_enc.p = 0.0
        m.preFit(BOWDataset(X, vocab))
        return m

    a = make()
    for _ in range(2):
        a.train_mb_delta()
        a.current_batch_sample = next(a.train_loader_iter)
    ck = str(tmp_path / "ck.pt")
    save_training_checkpoint(ck, a, a.train_loader, extra={"note": "after 2 steps"})
    torch_state = torch.get_rng_state()
    batch = a.current_batch_sample
    a.train_mb_delta()
    want = {k: v.clone() for k, v in a.model.state_dict().items()}

    b = make()
    extra = load_training_checkpoint(ck, b, b.train_loader)
    assert extra["note"] == "after 2 steps"
    b.current_batch_sample = batch
    torch.set_rng_state(torch_state)
    b.train_mb_delta()
    for k, v in b.model.state_dict().items():
        torch.testing.assert_close(v, want[k], rtol=1e-6, atol=1e-7, msg=lambda m: f"{k}: {m}")


def test_jsonl_metrics_and_timer(tmp_path):
    from gfedntm_b200.utils.observability import JsonlMetrics, StepTimer, nvtx_range

    m = JsonlMetrics(str(tmp_path / "m.jsonl"))
    m.log(step=1, loss=2.5)
    m.close()
    rec = json.loads(open(str(tmp_path / "m.jsonl")).readline())
    assert rec["step"] == 1 and rec["loss"] == 2.5 and "ts" in rec
    t = StepTimer(enabled=True)
    with t.region("x"), nvtx_range("x"):
        pass
    assert isinstance(t.summary(), dict)
