"""End-to-end gRPC federation on CPU: 1 server + 2 clients on localhost (BASELINE.json config 1,
scaled down): vocabulary consensus, initial-model distribution, synchronous per-minibatch training,
stop protocol and the npz result layout."""
import os
import socket
import threading
import time
from concurrent import futures

import grpc
import numpy as np
import pytest


def _free_port_block(n=4):
    for _ in range(50):
        s = socket.socket()
        s.bind(("127.0.0.1", 0))
        base = s.getsockname()[1]
        s.close()
        ok = True
        for i in range(n):
            t = socket.socket()
            try:
                t.bind(("127.0.0.1", base + i))
            except OSError:
                ok = False
            finally:
                t.close()
        if ok and base + n < 65000:
            return base
    raise RuntimeError("no free port block")


def _run_federation(tmp_path, model_type, agg, n_iters, embeddings=False, num_epochs=2):
    from gfedntm_b200.federation.client import Client
    from gfedntm_b200.federation.server import FederatedServer
    from gfedntm_b200.protos import federated_pb2_grpc
    from gfedntm_b200.utils.auxiliary_functions import read_config_experiments
    from gfedntm_b200.utils.synthetic import counts_to_documents, generate_federated_corpus

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    params = read_config_experiments(os.path.join(root, "config", "dft_params.cf"))
    params.update(n_components=5, hidden_sizes=(16, 16), num_epochs=num_epochs, batch_size=32, verbose=False,
                  num_samples=2)
    if embeddings:
        params["contextual_size"] = 8
    topics, doc_topics, counts = generate_federated_corpus(vocab_size=150, n_topics=5, n_docs=70, n_nodes=2,
                                                           frozen_topics=1, nwords=(20, 30), seed=3)
    base = _free_port_block(4)
    grads = ["prior_mean", "prior_variance", "beta", "inf_net.input_layer.weight", "inf_net.input_layer.bias",
             "inf_net.f_mu.weight", "beta_batchnorm.running_mean", "beta_batchnorm.num_batches_tracked",
             "inf_net.adapt_bert.weight", "inf_net.adapt_bert.bias"]
    save_server = str(tmp_path / "server" / "global_model")
    os.makedirs(os.path.dirname(save_server), exist_ok=True)
    fed = FederatedServer(min_num_clients=2, model_params=params, model_type=model_type, max_iters=n_iters,
                          opts_client=[], save_server=save_server, logs_server=str(tmp_path / "server.log"),
                          grads_to_share=grads, client_server_addres="localhost", base_port=base, agg=agg,
                          wait_timeout=60)
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=8))
    federated_pb2_grpc.add_FederationServicer_to_server(fed, server)
    server.add_insecure_port(f"127.0.0.1:{base}")
    server.start()
    clients, errors = {}, []

    def run_client(cid):
        try:
            docs = counts_to_documents(counts[cid - 1])
            if embeddings:
                import pandas as pd

                rng = np.random.default_rng(cid)
                corpus = pd.DataFrame({"bow_text": [" ".join(d) for d in docs],
                                       "embeddings": [rng.normal(size=8).astype(np.float32) for _ in docs]})
                dtype = "real"
            else:
                corpus, dtype = docs, "synthetic"
            ch = grpc.insecure_channel(f"127.0.0.1:{base}")
            stub = federated_pb2_grpc.FederationStub(ch)
            clients[cid] = Client(id=cid, stub=stub, local_corpus=corpus, data_type=dtype, opts_server=[],
                                  save_client=str(tmp_path / f"client{cid}" / f"model_{cid}"),
                                  logs_client=str(tmp_path / f"client{cid}.log"), sleep_time=120, base_port=base,
                                  grads_to_share=grads, share="grad" if agg == "grad" else "param",
                                  backend="eager", device="cpu")
        except BaseException as ex:  # surfaced in the main thread
            errors.append(ex)

    for cid in (1, 2):
        os.makedirs(str(tmp_path / f"client{cid}"), exist_ok=True)
    threads = [threading.Thread(target=run_client, args=(cid,)) for cid in (1, 2)]
    for t in threads:
        t.start()
    assert fed.training_finished.wait(timeout=180), "federated training did not finish"
    for t in threads:
        t.join(timeout=60)
    server.stop(grace=0.5)
    if fed.training_error is not None:
        raise fed.training_error
    assert not errors, errors
    return fed, clients, tmp_path


@pytest.mark.parametrize("agg", ["param", "grad"])
def test_avitm_federation_over_grpc(tmp_path, agg):
    fed, clients, out = _run_federation(tmp_path, "avitm", agg, n_iters=8)
    assert len(fed.iter_times) >= 6
    # every client ends with the same values for the shared tensors, equal to the server's global model
    g = fed._global_model.model.state_dict()
    for cid, c in clients.items():
        sd = c.local_model.model.state_dict()
        for k in ("beta", "prior_mean", "inf_net.input_layer.weight", "beta_batchnorm.running_mean"):
            np.testing.assert_allclose(sd[k].numpy(), g[k].numpy(), rtol=1e-5, atol=1e-6, err_msg=f"client {cid} {k}")
        # unshared tensors stay client-local
    a, b = (clients[i].local_model.model.state_dict()["inf_net.f_sigma.weight"].numpy() for i in (1, 2))
    assert not np.allclose(a, b)
    # result files in the reference layout
    for cid in (1, 2):
        z = np.load(str(out / f"client{cid}" / f"model_{cid}.npz"), allow_pickle=True)
        assert set(z.files) == {"betas", "thetas", "ntopics", "topics"}
        assert z["betas"].shape[0] == 5 and z["thetas"].shape == (70, 5) and z["topics"].shape == (5, 10)
        np.testing.assert_allclose(z["betas"].sum(1), 1.0, rtol=1e-4)
        np.testing.assert_allclose(z["thetas"].sum(1), 1.0, rtol=1e-6)
    zs = np.load(str(out / "server" / "global_model.npz"), allow_pickle=True)
    assert zs["betas"].shape[0] == 5 and int(zs["ntopics"]) == 5


def test_ctm_federation_stops_at_num_epochs(tmp_path):
    # 70 docs / batch 32 -> 3 minibatches per epoch, 1 epoch: clients finish long before max_iters
    fed, clients, out = _run_federation(tmp_path, "ctm", "param", n_iters=50, embeddings=True, num_epochs=1)
    assert len(fed.iter_times) < 10
    for cid in (1, 2):
        assert os.path.exists(str(out / f"client{cid}" / f"model_{cid}_epoch_1.npz"))  # CTM per-epoch snapshot
        assert os.path.exists(str(out / f"client{cid}" / f"model_{cid}.npz"))
        assert clients[cid].local_model.model.state_dict()["inf_net.adapt_bert.weight"].shape[1] == 8


def test_pull_is_idempotent_per_iteration_and_clients_checkpoint(tmp_path):
    """A retried ``getGradient`` for the same global iteration returns the cached payload instead of training a
    second time (server.py's ``_call`` retries pulls), and a client servicer writes / restores mid-training
    checkpoints."""
    import torch

    from gfedntm_b200.federated.federated_avitm import FederatedAVITM
    from gfedntm_b200.federation.client import FederatedClientServer
    from gfedntm_b200.models import BOWDataset
    from gfedntm_b200.protos import federated_pb2
    from gfedntm_b200.utils.auxiliary_functions import modelStateDict_to_proto

    def make(resume=False):
        torch.manual_seed(0)
        rng = np.random.default_rng(0)
        X = rng.poisson(0.3, (40, 60)).astype(np.float32)
        X[X.sum(1) == 0, 0] = 1
        params = dict(input_size=60, n_components=4, model_type="prodLDA", hidden_sizes=(8, 8), activation="softplus",
                      dropout=0.2, learn_priors=True, batch_size=16, lr=2e-3, momentum=0.99, solver="adam", num_epochs=5,
                      reduce_on_plateau=False, num_samples=2, verbose=False, seed=1, topic_prior_mean=0.0,
                      topic_prior_variance=None, num_data_loader_workers=0)
        import logging

        m = FederatedAVITM(params, ["beta", "prior_mean"], logging.getLogger("t"), backend="eager", device="cpu")
        srv = FederatedClientServer(m, BOWDataset(X, [f"w{i}" for i in range(60)]), 1, str(tmp_path / "model_1"),
                                    checkpoint_path=str(tmp_path / "client1.ckpt"), checkpoint_every=2, resume=resume)
        return m, srv

    m, srv = make()
    r1 = srv.getGradient(federated_pb2.ServerGetGradientRequest(iter=0), None)
    steps_after_first = int(m.optimizer.state_dict()["state"][0]["step"])
    r2 = srv.getGradient(federated_pb2.ServerGetGradientRequest(iter=0), None)
    assert r2 is r1 and int(m.optimizer.state_dict()["state"][0]["step"]) == steps_after_first
    # two aggregated updates -> a checkpoint exists; a fresh client restores the same weights and bookkeeping
    for it in range(2):
        if it:
            srv.getGradient(federated_pb2.ServerGetGradientRequest(iter=it), None)
        sd = {k: v for k, v in m.model.state_dict().items() if k in ("beta", "prior_mean")}
        req = federated_pb2.ServerAggregatedTensorRequest(
            header=federated_pb2.MessageHeader(message_type=federated_pb2.MessageType.SERVER_AGGREGATED_TENSOR_SEND),
            nndata=federated_pb2.NNUpdate(modelUpdate=modelStateDict_to_proto(sd, -1, "avitm")))
        ack = srv.sendAggregatedTensor(req, None)
        assert not ack.header.id_request.endswith(";finished")
    assert os.path.exists(tmp_path / "client1.ckpt")
    m2, srv2 = make(resume=True)
    assert m2.current_mb == m.current_mb and m2.samples_processed == m.samples_processed
    for k, v in m.model.state_dict().items():
        assert torch.equal(v, m2.model.state_dict()[k]), k
