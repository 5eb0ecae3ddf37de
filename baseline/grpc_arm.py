#!/usr/bin/env python
"""The reference's OWN gRPC federation loop as a measured baseline (variant (a) / (b)).

Runs the unmodified sources from ``baseline/_ref`` as one server process and N client processes on localhost
(``FederatedServer`` / ``Client`` constructed exactly like the reference's ``main.py:27-175`` does): vocabulary
consensus, initial-model distribution, then ``do_federated_training`` -- sequential ``getGradient`` pulls, the numpy
weighted mean and sequential ``sendAggregatedTensor`` pushes of ``src/federation/server.py:430-521``.  The per-iteration
time is read from the server's own log ("ACK received for iter i" time stamps).

No reference file is edited.  What is done from the outside:
  * environment shims of baseline/ref_arm.py (numpy.Inf, matplotlib stub) and a polling stand-in for the missing
    ``waiting`` package;
  * client address prefix ``127.0.0.`` so that the reference's ``<prefix><id>:<port>`` rule yields loopback addresses;
  * model_type ``ctm`` (CombinedTM): the reference's avitm server path raises at server.py:294;
  * ``--sleep 0``: the server module's ``time`` name is replaced by a proxy whose ``sleep`` returns immediately
    (variant (b), "gRPC, sleep removed"); ``--sleep ref`` keeps the hard-coded 3 s per client per step.
The run ends when the requested iterations are logged (the reference's server-side save crashes afterwards, 6.3 item 12).

    python baseline/grpc_arm.py --clients 2 --vocab 5000 --topics 50 --iters 8 --sleep 0
"""
from __future__ import annotations

import argparse
import json
import os
import re
import subprocess
import sys
import tempfile
import time
import types

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)


def _shims():
    from baseline.ref_arm import REF_ROOT, _shim_env, ensure_installed

    ensure_installed()
    _shim_env()
    w = types.ModuleType("waiting")

    def wait(predicate, timeout_seconds=120, sleep_seconds=0.05, **_):
        t0 = time.time()
        while not predicate():
            if time.time() - t0 > timeout_seconds:
                raise TimeoutError("waiting.wait timed out")
            time.sleep(sleep_seconds)
        return True

    w.wait = wait
    sys.modules["waiting"] = w
    if REF_ROOT not in sys.path:
        sys.path.insert(0, REF_ROOT)
    return REF_ROOT


def _params(ref_root, a):
    from src.utils.auxiliary_functions import read_config_experiments

    p = read_config_experiments(os.path.join(ref_root, "config", "dft_params.cf"))
    p.update(n_components=a.topics, batch_size=a.batch, num_epochs=1000, contextual_size=a.ctx, hidden_sizes=(50, 50),
             num_samples=2, verbose=False)
    return p


GRPC_OPTS = [("grpc.max_send_message_length", 1 << 30), ("grpc.max_receive_message_length", 1 << 30),
             ("grpc.max_message_length", 1 << 30)]


def role_server(a):
    ref_root = _shims()
    from concurrent import futures

    import grpc
    from src.federation import server as srv_mod
    from src.protos import federated_pb2_grpc

    if a.sleep == "0":
        class _T:  # everything of `time` except sleeping
            def __getattr__(self, k):
                return getattr(time, k)

            @staticmethod
            def sleep(_s):
                return None

        srv_mod.time = _T()
    grads = [g for g in _params(ref_root, a)["grads_to_share"]] if False else None
    import configparser

    cf = configparser.ConfigParser()
    cf.read(os.path.join(ref_root, "config", "dft_params.cf"))
    grads = cf.get("federation", "grads_to_share").split(",")
    server = grpc.server(futures.ThreadPoolExecutor(), options=GRPC_OPTS)
    fed = srv_mod.FederatedServer(min_num_clients=a.clients, model_params=_params(ref_root, a), model_type="ctm",
                                  max_iters=a.iters, opts_client=GRPC_OPTS, save_server=os.path.join(a.work, "global_model"),
                                  logs_server=os.path.join(a.work, "server.log"), grads_to_share=grads,
                                  client_server_addres="127.0.0.", base_port=a.port)
    federated_pb2_grpc.add_FederationServicer_to_server(fed, server)
    server.add_insecure_port(f"127.0.0.1:{a.port}")
    server.start()
    server.wait_for_termination(a.timeout)


def role_client(a):
    ref_root = _shims()
    import grpc
    import pandas as pd
    from src.federation.client import Client
    from src.protos import federated_pb2_grpc
    import configparser

    cf = configparser.ConfigParser()
    cf.read(os.path.join(ref_root, "config", "dft_params.cf"))
    grads = cf.get("federation", "grads_to_share").split(",")
    df = pd.read_parquet(os.path.join(a.work, f"client{a.id}.parquet"))
    os.makedirs(os.path.join(a.work, f"client{a.id}"), exist_ok=True)
    with grpc.insecure_channel(f"127.0.0.1:{a.port}", options=GRPC_OPTS) as ch:
        stub = federated_pb2_grpc.FederationStub(ch)
        Client(id=a.id, stub=stub, local_corpus=df, data_type="real", opts_server=GRPC_OPTS,
               save_client=os.path.join(a.work, f"client{a.id}", "model"), logs_client=os.path.join(a.work, f"client{a.id}.log"),
               sleep_time=a.timeout, base_port=a.port, grads_to_share=grads)


def role_server_ours(a):
    """Same launcher, this repository's gRPC control/data plane (gfedntm_b200.federation) for the side-by-side row."""
    from concurrent import futures

    import grpc

    from gfedntm_b200.federation.server import FederatedServer
    from gfedntm_b200.protos import federated_pb2_grpc
    from gfedntm_b200.utils.auxiliary_functions import read_config_experiments
    import configparser

    cfg = os.path.join(ROOT, "config", "dft_params.cf")
    p = read_config_experiments(cfg)
    p.update(n_components=a.topics, batch_size=a.batch, num_epochs=1000, contextual_size=a.ctx, hidden_sizes=(50, 50),
             num_samples=2, verbose=False)
    cf = configparser.ConfigParser()
    cf.read(cfg)
    grads = cf.get("federation", "grads_to_share").split(",")
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=16), options=GRPC_OPTS)
    fed = FederatedServer(min_num_clients=a.clients, model_params=p, model_type="ctm", max_iters=a.iters,
                          opts_client=GRPC_OPTS, save_server=os.path.join(a.work, "global_model"),
                          logs_server=os.path.join(a.work, "server.log"), grads_to_share=grads,
                          client_server_addres="127.0.0.", base_port=a.port, agg="param")
    federated_pb2_grpc.add_FederationServicer_to_server(fed, server)
    server.add_insecure_port(f"127.0.0.1:{a.port}")
    server.start()
    fed.training_finished.wait(a.timeout)
    time.sleep(1.0)
    server.stop(grace=0.5)


def role_client_ours(a):
    import grpc
    import pandas as pd

    from gfedntm_b200.federation.client import Client
    from gfedntm_b200.protos import federated_pb2_grpc
    import configparser

    cf = configparser.ConfigParser()
    cf.read(os.path.join(ROOT, "config", "dft_params.cf"))
    grads = cf.get("federation", "grads_to_share").split(",")
    df = pd.read_parquet(os.path.join(a.work, f"client{a.id}.parquet"))
    os.makedirs(os.path.join(a.work, f"client{a.id}"), exist_ok=True)
    with grpc.insecure_channel(f"127.0.0.1:{a.port}", options=GRPC_OPTS) as ch:
        stub = federated_pb2_grpc.FederationStub(ch)
        Client(id=a.id, stub=stub, local_corpus=df, data_type="real", opts_server=GRPC_OPTS,
               save_client=os.path.join(a.work, f"client{a.id}", "model"), logs_client=os.path.join(a.work, f"client{a.id}.log"),
               sleep_time=a.timeout, base_port=a.port, grads_to_share=grads, share="param", backend="auto")


def role_launch(a):
    import pandas as pd
    import torch

    from gfedntm_b200.utils.synthetic import generate_federated_corpus

    work = a.work or tempfile.mkdtemp(prefix="gfedntm_grpc_arm_")
    os.makedirs(work, exist_ok=True)
    topics, doc_topics, counts = generate_federated_corpus(vocab_size=a.vocab, n_topics=a.topics, n_docs=a.docs,
                                                           n_nodes=a.clients, frozen_topics=3, nwords=(150, 250), seed=5)
    rng = np.random.default_rng(0)
    for c in range(a.clients):
        m = counts[c].tocsr()
        texts = []
        for r in range(m.shape[0]):
            lo, hi = m.indptr[r], m.indptr[r + 1]
            texts.append(" ".join(f"wd{v}" for v, n in zip(m.indices[lo:hi], m.data[lo:hi]) for _ in range(int(n))))
        emb = rng.standard_normal((m.shape[0], a.ctx)).astype(np.float32)
        pd.DataFrame({"bow_text": texts, "fos": "x", "embeddings": [" ".join(f"{x:.5f}" for x in e) for e in emb]}
                     ).to_parquet(os.path.join(work, f"client{c + 1}.parquet"))
    n_gpu = torch.cuda.device_count()
    base = [sys.executable, os.path.abspath(__file__), "--work", work, "--port", str(a.port), "--clients", str(a.clients),
            "--vocab", str(a.vocab), "--topics", str(a.topics), "--ctx", str(a.ctx), "--batch", str(a.batch),
            "--iters", str(a.iters), "--sleep", a.sleep, "--timeout", str(a.timeout), "--impl", a.impl]
    env = dict(os.environ, PYTHONPATH=ROOT)
    procs = [subprocess.Popen(base + ["--role", "server"], env=dict(env, CUDA_VISIBLE_DEVICES=""),
                              stdout=open(os.path.join(work, "server.out"), "w"), stderr=subprocess.STDOUT)]
    time.sleep(4.0)
    for c in range(a.clients):
        e = dict(env, CUDA_VISIBLE_DEVICES=str(c % n_gpu) if n_gpu else "")
        procs.append(subprocess.Popen(base + ["--role", "client", "--id", str(c + 1)], env=e,
                                      stdout=open(os.path.join(work, f"client{c + 1}.out"), "w"), stderr=subprocess.STDOUT))
    log = os.path.join(work, "server.log")
    pat = re.compile(r"^(\d{4}-\d\d-\d\d \d\d:\d\d:\d\d),(\d{3}).*" +
                     (r"ACK received for iter (\d+)" if a.impl == "reference" else r"global iteration (\d+) done"))
    t_end = time.time() + a.timeout
    stamps = {}
    try:
        while time.time() < t_end:
            time.sleep(1.0)
            if os.path.exists(log):
                for ln in open(log, errors="ignore"):
                    mt = pat.match(ln)
                    if mt:
                        ts = time.mktime(time.strptime(mt.group(1), "%Y-%m-%d %H:%M:%S")) + int(mt.group(2)) / 1e3
                        stamps[int(mt.group(3))] = max(stamps.get(int(mt.group(3)), 0.0), ts)
            if len(stamps) >= a.iters or any(p.poll() not in (None, 0) for p in procs[1:]):
                break
    finally:
        for p in procs:
            if p.poll() is None:
                p.terminate()
        for p in procs:
            try:
                p.wait(timeout=10)
            except Exception:
                p.kill()
    its = sorted(stamps)
    per = [stamps[i] - stamps[i - 1] for i in its if i - 1 in stamps]
    V_eff = None
    out = {"impl": "reference gRPC loop (unmodified sources)" if a.impl == "reference" else
           "gfedntm_b200 gRPC federation (concurrent pulls, persistent channels, fp32 payloads)", "sleep": a.sleep, "clients": a.clients, "model": "CombinedTM",
           "vocab_requested": a.vocab, "topics": a.topics, "contextual_size": a.ctx, "batch_per_client": a.batch,
           "iterations_logged": len(its), "s_per_iteration": per,
           "median_s_per_iteration": float(np.median(per)) if per else None,
           "docs_per_s": (a.clients * a.batch / float(np.median(per))) if per else None, "work_dir": work}
    if not per:
        for f in ("server.out", "client1.out"):
            pth = os.path.join(work, f)
            if os.path.exists(pth):
                out[f] = open(pth, errors="ignore").read()[-1500:]
    print(json.dumps(out))
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--role", default="launch", choices=["launch", "server", "client"])
    ap.add_argument("--id", type=int, default=1)
    ap.add_argument("--work", default=None)
    ap.add_argument("--port", type=int, default=52051)
    ap.add_argument("--clients", type=int, default=2)
    ap.add_argument("--vocab", type=int, default=5000)
    ap.add_argument("--topics", type=int, default=50)
    ap.add_argument("--ctx", type=int, default=768)
    ap.add_argument("--batch", type=int, default=64)
    ap.add_argument("--docs", type=int, default=1280, help="documents per client (an epoch end triggers a snapshot)")
    ap.add_argument("--iters", type=int, default=8)
    ap.add_argument("--sleep", default="0", choices=["0", "ref"])
    ap.add_argument("--timeout", type=int, default=240)
    ap.add_argument("--impl", default="reference", choices=["reference", "ours"])
    a = ap.parse_args()
    roles = {"launch": role_launch, "server": role_server if a.impl == "reference" else role_server_ours,
             "client": role_client if a.impl == "reference" else role_client_ours}
    roles[a.role](a)


if __name__ == "__main__":
    main()
