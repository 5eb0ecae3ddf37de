"""CPU tests of the framework layer: networks, trainers, datasets, loader, serialisation, config,
metrics, synthetic generator, registry."""
import configparser
import os
import sys

import numpy as np
import pytest
import scipy.sparse as sp
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _toy(n=120, v=200, seed=0):
    rng = np.random.default_rng(seed)
    X = rng.poisson(0.08, (n, v)).astype(np.float32)
    X[X.sum(1) == 0, 0] = 1
    return X, [f"w{i}" for i in range(v)]


def test_state_dict_keys_match_reference_layout():
    from gfedntm_b200.models.networks import DecoderNetwork

    m = DecoderNetwork(2000, 50, "prodLDA", (50, 50), "softplus", 0.2, True)
    sd = m.state_dict()
    assert len(sd) == 20 and sum(p.numel() for p in m.parameters()) == 207800
    for k in ("prior_mean", "prior_variance", "beta", "inf_net.input_layer.weight", "inf_net.hiddens.l_0.0.weight",
              "inf_net.f_mu_batchnorm.running_mean", "inf_net.f_sigma_batchnorm.num_batches_tracked",
              "beta_batchnorm.running_var"):
        assert k in sd
    c = DecoderNetwork(2000, 50, "prodLDA", (50, 50), "softplus", 0.2, True, infnet="combined", contextual_size=768)
    assert sum(p.numel() for p in c.parameters()) == 1845800
    assert c.state_dict()["inf_net.adapt_bert.weight"].shape == (2000, 768)
    assert c.state_dict()["inf_net.input_layer.weight"].shape == (50, 4000)
    m(torch.rand(4, 2000))
    assert "topic_word_matrix" not in m.state_dict()


@pytest.mark.parametrize("act", ["softplus", "relu", "sigmoid", "tanh", "leakyrelu", "rrelu", "elu", "selu"])
def test_all_activations_forward_backward(act):
    from gfedntm_b200.models.networks import DecoderNetwork, elbo_terms

    m = DecoderNetwork(60, 5, "LDA", (16,), act, 0.2, True)
    x = torch.rand(8, 60)
    pm, pv, mu, var, lv, wd = m(x)
    kl, rl = elbo_terms(x, wd, pm, pv, mu, var, lv)
    (kl + rl).sum().backward()
    assert torch.isfinite(m.beta.grad).all()
    m.eval()  # without theta-dropout the LDA word distribution is a proper distribution
    torch.testing.assert_close(m(x)[5].sum(1), torch.ones(8), rtol=1e-4, atol=1e-4)


def test_forward_matches_reference_modules():
    """Parity with the reference's own DecoderNetwork / _loss on CPU (same weights, eval-free path)."""
    ref_root = "/root/reference"
    if not os.path.isdir(ref_root):
        pytest.skip("reference tree not mounted")
    np.Inf = np.inf  # NumPy 2 shim needed by the vendored early-stopping helper
    sys.path.insert(0, ref_root)
    try:
        from src.models.base.pytorchavitm.avitm_network.decoder_network import DecoderNetwork as RefNet
    finally:
        sys.path.remove(ref_root)
    from gfedntm_b200.models.networks import DecoderNetwork, elbo_terms

    torch.manual_seed(0)
    for mt in ("prodLDA", "LDA"):
        ours = DecoderNetwork(300, 7, mt, (20, 20), "softplus", 0.0, True)
        ref = RefNet(300, 7, mt, (20, 20), "softplus", 0.0, True)
        ref.load_state_dict(ours.state_dict())
        ours.inf_net.dropout_enc.p = 0.0
        ref.inf_net.dropout_enc.p = 0.0
        x = torch.poisson(torch.full((16, 300), 0.1))
        torch.manual_seed(5)
        o = ours(x)
        torch.manual_seed(5)
        r = ref(x)
        for a, b in zip(o, r):
            torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6)
        kl, rl = elbo_terms(x, o[5], o[0], o[1], o[2], o[3], o[4])
        var_division = torch.sum(r[3] / r[1], dim=1)
        diff = r[0] - r[2]
        ref_kl = 0.5 * (var_division + torch.sum(diff * diff / r[1], dim=1) - 7 + r[1].log().sum() - r[4].sum(dim=1))
        torch.testing.assert_close(kl, ref_kl, rtol=1e-5, atol=1e-5)


def test_avitm_fit_save_load_and_results(tmp_path):
    from gfedntm_b200.models import AVITM, BOWDataset

    X, vocab = _toy()
    ds = BOWDataset(X, vocab)
    m = AVITM(input_size=200, n_components=4, hidden_sizes=(16, 16), num_epochs=3, verbose=False, batch_size=32,
              device="cpu", seed=1)
    assert m.backend == "eager"
    m.fit(ds, validation_dataset=BOWDataset(X[:40], vocab), save_dir=str(tmp_path))
    assert m.training_doc_topic_distributions.shape == (120, 4)
    np.testing.assert_allclose(m.training_doc_topic_distributions.sum(1), 1.0, rtol=1e-4)
    topics = m.get_topics(5)
    assert len(topics) == 4 and all(len(t) == 5 and t[0].startswith("w") for t in topics)
    assert m.get_topic_word_distribution().shape == (4, 200)
    assert len(m.get_predicted_topics(ds, 2)) == 120
    m.nn_epoch = 0
    m.save(str(tmp_path))
    path = os.path.join(str(tmp_path), m._format_file())
    assert os.path.exists(os.path.join(path, "epoch_0.pth"))
    m2 = AVITM(input_size=200, n_components=4, hidden_sizes=(16, 16), num_epochs=3, verbose=False, device="cpu")
    m2.load(path, 0)
    torch.testing.assert_close(m2.model.beta, m.model.beta)


@pytest.mark.parametrize("kind", ["combined", "zeroshot"])
def test_ctm_fit(kind):
    from gfedntm_b200.models import CTM, CTMDataset

    X, vocab = _toy()
    ctx = np.random.default_rng(1).normal(size=(120, 12)).astype(np.float32)
    ds = CTMDataset(ctx, X, vocab)
    m = CTM(input_size=200, contextual_size=12, inference_type=kind, n_components=3, hidden_sizes=(8,),
            num_epochs=2, verbose=False, device="cpu", batch_size=40)
    m.fit(ds)
    assert np.isfinite(m.best_loss_train)
    assert m.get_doc_topic_distribution(ds, 2).shape == (120, 3)
    assert len(m.get_word_distribution_by_topic_id(0)) == 200


def test_labels_path_runs_on_eager():
    from gfedntm_b200.models import CTM, CTMDataset

    X, vocab = _toy(60, 80)
    ctx = np.random.default_rng(2).normal(size=(60, 6)).astype(np.float32)
    labels = np.eye(3, dtype=np.float32)[np.arange(60) % 3]
    m = CTM(input_size=80, contextual_size=6, n_components=3, hidden_sizes=(8,), num_epochs=1, verbose=False,
            device="cpu", label_size=3, batch_size=20)
    m.fit(CTMDataset(ctx, X, vocab, labels=labels))
    assert "label_classification.weight" in m.model.state_dict()


def test_loader_batches_cover_corpus_and_match_dense():
    from gfedntm_b200.models import BOWDataset
    from gfedntm_b200.utils.loader import CsrMinibatchLoader

    X, vocab = _toy(101, 50)
    ds = BOWDataset(X, vocab)
    seen = []
    for mb in CsrMinibatchLoader(ds, 32, shuffle=True, seed=3):
        dense = mb.bow.to_dense().numpy()
        np.testing.assert_array_equal(dense, X[mb.index])
        seen.extend(mb.index.tolist())
    assert sorted(seen) == list(range(101))
    assert len(CsrMinibatchLoader(ds, 32, drop_last=True)) == 3
    np.testing.assert_array_equal(ds[5]["X"].numpy(), X[5])
    # endless iteration: every epoch is a permutation of the corpus, the iterator never ends
    it = CsrMinibatchLoader(ds, 32, shuffle=True, seed=3).cycle()
    epochs = [[], []]
    for b in range(8):           # 4 batches per epoch (32 + 32 + 32 + 5)
        mb = next(it)
        np.testing.assert_array_equal(mb.bow.to_dense().numpy(), X[mb.index])
        epochs[b // 4].extend(mb.index.tolist())
    it.close()
    assert sorted(epochs[0]) == list(range(101)) and sorted(epochs[1]) == list(range(101))
    assert epochs[0] != epochs[1]


def test_tensor_and_model_serialisation_roundtrip():
    from gfedntm_b200.models.networks import DecoderNetwork
    from gfedntm_b200.utils import auxiliary_functions as aux

    t = torch.randn(3, 5)
    torch.testing.assert_close(aux.deserializeTensor(aux.serializeTensor(t)), t)
    b = torch.randn(4, 4).to(torch.bfloat16)
    torch.testing.assert_close(aux.deserializeTensor(aux.serializeTensor(b)), b.float())
    i = torch.tensor(7, dtype=torch.long)
    assert int(aux.deserializeTensor(aux.serializeTensor(i))) == 7
    m = DecoderNetwork(40, 4, "prodLDA", (8, 8, 8), "relu", 0.2, True)  # 3 hidden layers: extra_tensors path
    sd = m.state_dict()
    back = aux.proto_to_modelStateDict(aux.modelStateDict_to_proto(sd, -1, "avitm"))
    assert set(back) == set(sd)
    for k in sd:
        torch.testing.assert_close(back[k].to(sd[k].dtype).reshape(sd[k].shape), sd[k])
    m2 = DecoderNetwork(40, 4, "prodLDA", (8, 8, 8), "relu", 0.2, True)
    m2.load_state_dict(back)


def test_reference_can_parse_our_model_update():
    ref_root = "/root/reference"
    if not os.path.isdir(ref_root):
        pytest.skip("reference tree not mounted")
    sys.path.insert(0, ref_root)
    try:
        from src.protos import federated_pb2 as ref_pb
    finally:
        sys.path.remove(ref_root)
    from gfedntm_b200.models.networks import DecoderNetwork
    from gfedntm_b200.utils import auxiliary_functions as aux

    m = DecoderNetwork(30, 3, "prodLDA", (6, 6), "softplus", 0.2, True)
    wire = aux.modelStateDict_to_proto(m.state_dict(), -1, "avitm").SerializeToString()
    ref = ref_pb.ModelUpdate.FromString(wire)
    assert ref.HasField("beta") and ref.beta.dtype == "float32"
    got = np.frombuffer(ref.beta.tensor_content, dtype=np.float32).reshape(3, 30)
    np.testing.assert_array_equal(got, m.beta.detach().numpy())
    assert ref.HasField("inf_net_hiddens_l00_weight") and ref.HasField("inf_net_hiddens_l_00_bias")


def test_optimizer_state_roundtrip():
    from gfedntm_b200.utils import auxiliary_functions as aux

    p = torch.nn.Parameter(torch.randn(4, 3))
    opt = torch.optim.Adam([p], lr=2e-3, betas=(0.99, 0.99))
    p.grad = torch.randn(4, 3)
    opt.step()
    sd = aux.proto_to_optStateDict(aux.optStateDict_to_proto(opt.state_dict()))
    torch.testing.assert_close(sd["state"][0]["exp_avg"], opt.state_dict()["state"][0]["exp_avg"])
    assert abs(sd["param_groups"][0]["lr"] - 2e-3) < 1e-9 and sd["param_groups"][0]["betas"][0] == pytest.approx(0.99)


def test_config_reader_matches_reference_types():
    from gfedntm_b200.utils.auxiliary_functions import read_config_experiments

    cfg = read_config_experiments(os.path.join(ROOT, "config", "dft_params.cf"))
    assert cfg["n_components"] == 50 and cfg["hidden_sizes"] == (50, 50) and cfg["learn_priors"] is True
    assert cfg["topic_prior_variance"] is None and cfg["labels"] == "" and cfg["lr"] == pytest.approx(0.002)
    assert cfg["contextual_size"] == 768 and cfg["model_type"] == "prodLDA"
    ref_cfg = "/root/reference/config/dft_params.cf"
    if os.path.exists(ref_cfg):
        ref = read_config_experiments(ref_cfg)
        for k in ("n_components", "batch_size", "hidden_sizes", "momentum", "num_epochs", "thetas_thr", "solver"):
            assert ref[k] == cfg[k]
        c = configparser.ConfigParser()
        c.read(ref_cfg)
        ours = configparser.ConfigParser()
        ours.read(os.path.join(ROOT, "config", "dft_params.cf"))
        assert ours.get("federation", "grads_to_share") == c.get("federation", "grads_to_share")


def test_metrics_and_synthetic_generator(tmp_path):
    from gfedntm_b200.utils import metrics
    from gfedntm_b200.utils.synthetic import generate_federated_corpus, save_synthetic_npz

    topics, doc_topics, counts = generate_federated_corpus(vocab_size=300, n_topics=10, n_docs=40, n_nodes=3,
                                                           frozen_topics=2, nwords=(30, 40), seed=0)
    assert topics.shape == (10, 300) and len(counts) == 3 and counts[0].shape == (40, 300)
    np.testing.assert_allclose(topics.sum(1), 1.0, rtol=1e-6)
    assert metrics.tss(topics, topics) == pytest.approx(10.0, rel=1e-6)
    assert metrics.dss(doc_topics[0], doc_topics[0]) == pytest.approx(0.0, abs=1e-9)
    perm = np.random.default_rng(0).permutation(10)
    assert metrics.tss(topics[perm], topics) == pytest.approx(10.0, rel=1e-6)
    assert 0 < metrics.topic_diversity(topics, 10) <= 1
    assert -1 <= metrics.npmi_coherence(topics, sp.vstack(counts), 5) <= 1
    assert 0 <= metrics.inverted_rbo(topics, 5) <= 1
    t = metrics.thresholded_thetas(np.array([[0.001, 0.5, 0.499]]))
    assert t[0, 0] == 0 and t.sum() == pytest.approx(1.0)
    path = str(tmp_path / "synthetic.npz")
    save_synthetic_npz(path, topics, doc_topics, counts)
    data = np.load(path, allow_pickle=True)
    assert int(data["vocab_size"]) == 300 and len(data["documents"]) == 3
    assert data["documents"][0][0][0].startswith("wd")


def test_federation_registry():
    from gfedntm_b200.federation import Federation, FederationClient

    f = Federation()
    f.connect_consensus("ipv4:1")
    f.connect_consensus("ipv4:2")
    FederationClient.set_id_by_key("ipv4:1", f.federation_clients, 1)
    FederationClient.set_id_by_key("ipv4:2", f.federation_clients, 2)
    FederationClient.set_can_start_training("ipv4:2", f.federation_clients, True)
    assert f.getClients() == ["ipv4:1", "ipv4:2"]
    assert f.client_by_id(2).ready_for_training and not f.client_by_id(1).ready_for_training
    f.disconnect("ipv4:1")
    assert len(f.federation_clients) == 2  # the roster is kept
    f.connect_update("ipv4:1", {"beta": np.zeros(2)}, 3, 1, "ID1_0", 10)
    c = f.client_by_id(1)
    assert c.current_mb == 3 and c.current_epoch == 1 and c.num_max_epochs == 10


def test_weighted_aggregation_matches_reference_formula():
    from gfedntm_b200.federation.server import aggregate_weighted

    rng = np.random.default_rng(0)
    tensors = [{"beta": rng.normal(size=(3, 4)).astype(np.float32), "n": np.array(5 + i, dtype=np.int64)}
               for i in range(3)]
    nr = [10, 20, 70]
    got = aggregate_weighted(tensors, nr, ["beta", "n"])
    # numpy restatement of server.py:477-487
    N = np.array(nr)
    want = np.sum(np.stack([t["beta"] * n for t, n in zip(tensors, nr)]), axis=0) / N.sum()
    np.testing.assert_allclose(got["beta"], want, rtol=1e-6)
    assert got["beta"].dtype == np.float32 and got["n"].dtype == np.int64


def test_resource_sampler_writes_rows(tmp_path):
    import csv
    import time

    from gfedntm_b200.utils.mem_usage import ResourceSampler

    path = tmp_path / "usage.csv"
    with ResourceSampler(str(path), period_s=0.05) as rs:
        time.sleep(0.2)
    rows = list(csv.reader(open(path)))
    assert rows[0][:6] == ResourceSampler.HEADER and len(rows) >= 3 and rs.rows_written == len(rows) - 1
    assert int(rows[1][3]) > 0  # rss


def test_hierarchical_corpus_selection():
    import numpy as np
    import scipy.sparse as sp

    from gfedntm_b200.aux_modules.hierarchical import htm_ds, htm_ws

    rng = np.random.default_rng(0)
    K, V, D = 3, 12, 40
    betas = np.zeros((K, V))
    for k in range(K):
        betas[k, 4 * k:4 * k + 4] = 0.25  # disjoint vocabularies per topic
    thetas = rng.dirichlet(np.full(K, 0.3), D)
    counts = sp.csr_matrix(rng.poisson(1.0, (D, V)))
    sub, keep = htm_ds(counts, thetas, topic=1, thr=0.5)
    assert (thetas[keep, 1] >= 0.5).all() and sub.shape[0] == keep.size
    ws, keep_ws = htm_ws(counts, thetas, betas, topic=2)
    dense = ws.toarray()
    assert dense[:, :8].sum() == 0  # only words of topic 2's vocabulary survive the arg-max selection
    np.testing.assert_array_equal(dense[:, 8:], counts[keep_ws].toarray()[:, 8:] * (thetas[keep_ws, 2:3] > 0))
    ws2, _ = htm_ws(counts, thetas, betas, topic=2, sample=True, seed=1)
    assert ws2.toarray()[:, :8].sum() == 0


def test_public_signatures_cover_the_reference():
    """Every constructor argument and public method of the reference's user-facing classes exists here."""
    import importlib
    import inspect
    import types

    ref_root = "/root/reference"
    if not os.path.isdir(ref_root):
        pytest.skip("reference tree not mounted")
    np.Inf = np.inf
    for name in ("matplotlib", "matplotlib.pyplot", "waiting", "wordcloud"):
        if name not in sys.modules:
            try:
                __import__(name)
            except Exception:
                m = types.ModuleType(name)
                m.wait = lambda *a, **k: True
                sys.modules[name] = m
    sys.path.insert(0, ref_root)
    pairs = [
        ("src.models.base.pytorchavitm.avitm_network.avitm", "AVITM", "gfedntm_b200.models", "AVITM"),
        ("src.models.base.contextualized_topic_models.ctm_network.ctm", "CTM", "gfedntm_b200.models", "CTM"),
        ("src.models.base.contextualized_topic_models.ctm_network.ctm", "ZeroShotTM", "gfedntm_b200.models", "ZeroShotTM"),
        ("src.models.base.contextualized_topic_models.ctm_network.ctm", "CombinedTM", "gfedntm_b200.models", "CombinedTM"),
        ("src.models.base.pytorchavitm.datasets.bow_dataset", "BOWDataset", "gfedntm_b200.models", "BOWDataset"),
        ("src.models.base.contextualized_topic_models.datasets.dataset", "CTMDataset", "gfedntm_b200.models", "CTMDataset"),
        ("src.models.federated.federated_model", "FederatedModel", "gfedntm_b200.federated.federated_model", "FederatedModel"),
        ("src.models.federated.federated_avitm", "FederatedAVITM", "gfedntm_b200.federated.federated_avitm", "FederatedAVITM"),
        ("src.models.federated.federated_ctm", "FederatedCTM", "gfedntm_b200.federated.federated_ctm", "FederatedCTM"),
        ("src.federation.server", "FederatedServer", "gfedntm_b200.federation.server", "FederatedServer"),
        ("src.federation.client", "Client", "gfedntm_b200.federation.client", "Client"),
        ("src.federation.client", "FederatedClientServer", "gfedntm_b200.federation.client", "FederatedClientServer"),
        ("src.federation.federation", "Federation", "gfedntm_b200.federation.federation", "Federation"),
        ("src.federation.federation_client", "FederationClient", "gfedntm_b200.federation.federation_client", "FederationClient"),
        ("src.models.base.utils.early_stopping.pytorchtools", "EarlyStopping", "gfedntm_b200.models.early_stopping", "EarlyStopping"),
    ]
    try:
        problems = []
        for rm, rn, om, on in pairs:
            ref = getattr(importlib.import_module(rm), rn)
            ours = getattr(importlib.import_module(om), on)
            pr = [p for p in inspect.signature(ref.__init__).parameters if p != "self"]
            po = [p for p in inspect.signature(ours.__init__).parameters if p != "self"]
            missing = [p for p in pr if p not in po]
            is_fn = lambda x: inspect.isfunction(x) or inspect.ismethod(x)  # noqa: E731
            rmeth = {n for n, _ in inspect.getmembers(ref, is_fn) if not n.startswith("__")}
            ometh = {n for n, _ in inspect.getmembers(ours, is_fn)}
            if missing or (rmeth - ometh):
                problems.append((rn, missing, sorted(rmeth - ometh)))
        import src.utils.auxiliary_functions as raf

        import gfedntm_b200.utils.auxiliary_functions as oaf

        rf = {n for n, v in inspect.getmembers(raf, inspect.isfunction) if v.__module__ == raf.__name__}
        of = {n for n, _ in inspect.getmembers(oaf, inspect.isfunction)}
        assert not problems and not (rf - of), (problems, sorted(rf - of))
        # module-level helpers of the side modules (skipped when a third-party import of the reference fails)
        for rm, om in (("src.aux_modules.utils.misc", "gfedntm_b200.aux_modules.misc"),
                       ("src.models.base.contextualized_topic_models.utils.data_preparation",
                        "gfedntm_b200.models.data_preparation"),
                       ("src.models.base.pytorchavitm.utils.data_preparation", "gfedntm_b200.models.data_preparation")):
            try:
                r = importlib.import_module(rm)
            except Exception:
                continue
            o = importlib.import_module(om)
            rfn = {n for n, v in inspect.getmembers(r, inspect.isfunction) if v.__module__ == r.__name__}
            ofn = {n for n, _ in inspect.getmembers(o, inspect.isfunction)}
            assert not (rfn - ofn), (rm, sorted(rfn - ofn))
    finally:
        sys.path.remove(ref_root)
        for k in [k for k in sys.modules if k == "src" or k.startswith("src.")]:
            del sys.modules[k]


def test_misc_keyboard_and_pickle_helpers(tmp_path):
    import warnings

    from gfedntm_b200.aux_modules import misc
    from gfedntm_b200.models.data_preparation import check_max_local_length

    ask = lambda ans: (lambda prompt: ans)  # noqa: E731
    assert misc.var_string_keyboard("comma_separated", (50, 50), "hidden", reader=ask("(100, 50)")) == (100, 50)
    assert misc.var_string_keyboard("comma_separated", (50, 50), "hidden", reader=ask("100")) == (50, 50)
    assert misc.var_string_keyboard("bool", True, "flag", reader=ask("False")) is False
    assert misc.var_string_keyboard("bool", True, "flag", reader=ask("nope")) is True
    assert misc.var_string_keyboard("dict", {"a": 1}, "d", reader=ask('{"b": 2}')) == {"b": 2}
    assert misc.var_string_keyboard("dict", {"a": 1}, "d", reader=ask("None")) == {"a": 1}
    assert misc.var_string_keyboard("str", "x", "s", reader=ask("")) == "x"
    assert misc.var_string_keyboard("str", "x", "s", reader=ask("abc")) == "abc"
    assert misc.format_title('A "quoted"\ntitle') == "A quotedtitle"
    path = str(tmp_path / "ob.pkl")
    misc.pickler(path, {"k": [1, 2, 3]})
    assert misc.unpickler(path) == {"k": [1, 2, 3]}
    with warnings.catch_warnings(record=True) as w:
        warnings.simplefilter("always")
        check_max_local_length(3, ["a b c d e", "a"])
        check_max_local_length(10, ["a b c"])
    assert len(w) == 1 and "5 words" in str(w[0].message)


def test_tmwrapper_hierarchical_submodels():
    from gfedntm_b200.aux_modules.tm_wrapper import TMWrapper

    rng = np.random.default_rng(0)
    vocab = [f"wd{i}" for i in range(60)]
    corpus = []
    for i in range(120):
        lo = (i % 3) * 20
        corpus.append(" ".join(rng.choice(vocab[lo:lo + 20], size=30)))
    tw = TMWrapper()
    father, ds, _ = tw.train_root_model(corpus, trainer="prod", ntopics=3, num_epochs=3, max_features=None,
                                        batch_size=32, device="cpu")
    child, cds, keep, secs = tw.train_htm_submodel("HTM-DS", father, ds, expansion_topic=0, thr=0.0, ntopics=2,
                                                   num_epochs=1, n_samples=2, batch_size=32, device="cpu")
    assert len(keep) == 120 and len(cds) == 120 and child.get_topic_word_distribution().shape == (2, 60)
    child2, cds2, keep2, _ = tw.train_htm_submodel("HTM-WS", father, ds, expansion_topic=1, ntopics=2, num_epochs=1,
                                                   n_samples=2, batch_size=16, device="cpu")
    assert 2 <= len(keep2) <= 120 and cds2.X.sum() <= ds.X.sum()


def test_tmmodel_folder_and_curation(tmp_path):
    from gfedntm_b200.aux_modules.tm_model import TMmodel

    rng = np.random.default_rng(0)
    K, V, D = 5, 40, 60
    betas = rng.dirichlet(np.full(V, 0.1), K)
    thetas = rng.dirichlet(np.full(K, 0.3), D)
    bow = sp.csr_matrix(rng.poisson(0.3, (D, V)))
    vocab = [f"wd{i}" for i in range(V)]
    tm = TMmodel(tmp_path / "model")
    tm.create(betas=betas, thetas=thetas, vocab=vocab, reference_bow=bow)
    a = tm.get_alphas()
    assert np.all(np.diff(a) <= 1e-12) and abs(a.sum() - 1.0) < 1e-6  # sorted by size, thetas row-normalised
    df = tm.to_dataframe()
    assert len(df) == K and (df["topic_entropy"] > 0).all() and (df["topic_entropy"] <= 1).all()
    assert np.isfinite(df["topic_coherence"]).all()
    for name in ("alphas.npy", "betas.npy", "thetas.npz", "vocab.txt", "betas_ds.npy", "tpc_descriptions.txt",
                 "alphas_orig.npy", "edits.txt"):
        assert (tmp_path / "model" / name).exists()
    # fuse: the merged topic carries the summed document weight
    before = tm.get_thetas().toarray()
    assert tm.fuseTopics([1, 3]) == 1
    after = tm.get_thetas().toarray()
    assert after.shape[1] == K - 1 and np.allclose(after[:, 1], before[:, 1] + before[:, 3])
    assert tm.deleteTopics([0]) == 1 and tm.get_betas().shape[0] == K - 2
    assert np.allclose(np.asarray(tm.get_thetas().sum(1)).ravel()[np.asarray(tm.get_thetas().sum(1)).ravel() > 0], 1.0)
    # a second object on the same folder sees the edited state; reset restores the original
    tm2 = TMmodel(tmp_path / "model")
    assert tm2.get_betas().shape[0] == K - 2 and tm2._edits == ["f 1 3", "d 0"]
    assert tm2.resetTM() == 1 and tm2.get_betas().shape[0] == K and tm2._edits == []
    sims = tm2.getSimilarTopics(3)
    assert len(sims["Coocurring"]) == 3 and 0.0 <= tm2.calculate_topic_diversity(10) <= 1.0


def test_inferencer_from_saved_checkpoint(tmp_path):
    from gfedntm_b200.aux_modules.inferencer import Inferencer
    from gfedntm_b200.models import AVITM, BOWDataset

    X, vocab = _toy(n=96, v=80, seed=3)
    m = AVITM(input_size=80, n_components=4, hidden_sizes=(16,), batch_size=32, num_epochs=2, verbose=False,
              device="cpu")
    m.fit(BOWDataset(X, vocab), save_dir=str(tmp_path))
    model_dir = [os.path.join(tmp_path, d) for d in os.listdir(tmp_path) if os.path.isdir(os.path.join(tmp_path, d))][0]
    inf = Inferencer(model_dir, device="cpu")
    assert inf.vocab == vocab
    texts = [" ".join(vocab[i] for i in np.flatnonzero(X[r])) for r in range(10)]
    th = inf.predict(texts, n_samples=3, out_dir=str(tmp_path / "inf"))
    assert th.shape == (10, 4) and np.allclose(th.sum(1), 1.0)
    assert np.load(tmp_path / "inf" / "doc-topics.npy").shape == (10, 4)
    # identical input through the document-term matrix gives the same distribution up to sampling noise
    th2 = inf.predict(bow=sp.csr_matrix((X[:10] > 0).astype(np.float32)), n_samples=3)
    assert th2.shape == (10, 4)


def test_prometheus_exporter_counts_iterations():
    from gfedntm_b200.utils.observability import PrometheusExporter

    off = PrometheusExporter(None)
    off.observe_iteration(0.1, 2)  # disabled: no-op
    assert off.scrape() == ""
    ex = PrometheusExporter(0)  # registry only, no HTTP server
    if not ex.enabled:
        pytest.skip("prometheus_client not installed")
    ex.observe_iteration(0.02, 3, payload_bytes=1000, failures=1)
    ex.observe_iteration(0.03, 2)
    text = ex.scrape()
    assert "gfedntm_global_iterations_total 2.0" in text and "gfedntm_active_clients 2.0" in text
    assert "gfedntm_payload_bytes_total 1000.0" in text and "gfedntm_rpc_failures_total 1.0" in text


def test_fabric_supervisor_restarts_from_checkpoint_and_gives_up():
    """Failure recovery of the fabric path: a launch that dies is repeated with resume=True while restarts are left."""
    from gfedntm_b200.federated.fabric_run import supervise

    calls = []

    def flaky(resume, attempt):
        calls.append((resume, attempt))
        if attempt < 2:
            raise RuntimeError("rank 3 exited with code -6")

    assert supervise(flaky, max_restarts=3, can_resume=True) == 2
    assert calls == [(False, 0), (True, 1), (True, 2)]
    calls.clear()
    with pytest.raises(RuntimeError):          # no checkpoints -> nothing to resume from: fail at once
        supervise(flaky, max_restarts=3, can_resume=False)
    assert calls == [(False, 0)]
    calls.clear()
    with pytest.raises(RuntimeError):          # restarts exhausted
        supervise(flaky, max_restarts=1, can_resume=True)
    assert calls == [(False, 0), (True, 1)]
