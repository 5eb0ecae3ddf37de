"""`main.py` as the user runs it: one server process and two client processes on localhost."""
import configparser
import os
import socket
import subprocess
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _free_base():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p if p + 3 < 65000 else p - 10


def test_main_server_and_two_clients(tmp_path):
    from gfedntm_b200.utils.synthetic import generate_federated_corpus, save_synthetic_npz

    topics, doc_topics, counts = generate_federated_corpus(vocab_size=120, n_topics=4, n_docs=40, n_nodes=2,
                                                           frozen_topics=1, nwords=(15, 25), seed=1)
    npz = str(tmp_path / "synthetic.npz")
    save_synthetic_npz(npz, topics, doc_topics, counts)
    base = _free_base()
    cfg = configparser.ConfigParser()
    cfg.read(os.path.join(ROOT, "config", "dft_params.cf"))
    cfg.set("addresses", "local", f"localhost:{base}")
    cfg.set("addresses", "base_port", str(base))
    cfg.set("federation", "server_port", str(base))
    cfg.set("federation", "time_termination", "120")
    cfg.set("federation", "client_sleep_time", "120")
    cfg.set("ntms", "n_components", "4")
    cfg.set("ntms", "hidden_sizes", "(8, 8)")
    cfg.set("ntms", "num_epochs", "1")
    cfg.set("ntms", "batch_size", "16")
    cfg.set("ntms", "num_samples", "2")
    cfg.set("b200", "addressing", "local")
    cfg.set("b200", "backend", "eager")
    for k in ("save_client", "save_server", "logs_client", "logs_server"):
        cfg.set("save_dir", k, str(tmp_path / "out" / k.replace("save_", "models_").replace("logs_", "logs_")))
    cfg_path = str(tmp_path / "params.cf")
    with open(cfg_path, "w") as f:
        cfg.write(f)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH=ROOT)
    common = [sys.executable, os.path.join(ROOT, "main.py"), "--config", cfg_path, "--source", npz, "--local"]
    server = subprocess.Popen(common + ["--id", "0", "--min_clients_federation", "2", "--max_iters", "6",
                                        "--model_type", "avitm"], env=env, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True)
    time.sleep(3.0)
    clients = [subprocess.Popen(common + ["--id", str(i), "--data_type", "synthetic"], env=env,
                                stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True) for i in (1, 2)]
    try:
        out_s, _ = server.communicate(timeout=170)
        outs = [c.communicate(timeout=60)[0] for c in clients]
    finally:
        for p in [server] + clients:
            if p.poll() is None:
                p.kill()
    assert server.returncode == 0, out_s[-3000:]
    for c, o in zip(clients, outs):
        assert c.returncode == 0, o[-3000:]
    stamp = time.strftime("%Y%m%d")
    base_out = tmp_path / "out"
    server_files = list((base_out / "models_server").glob("global_model_*.npz"))
    assert server_files, os.listdir(str(base_out))
    for cid in (1, 2):
        files = list(base_out.glob(f"models_client{cid}/model_{cid}_*.npz"))
        assert files, [str(p) for p in base_out.rglob("*")]
        z = np.load(str(files[0]), allow_pickle=True)
        assert z["betas"].shape[0] == 4 and z["thetas"].shape == (40, 4)


def test_main_real_data_ctm_through_the_sample_parquet(tmp_path):
    """`--data_type real --model_type ctm` end to end on the shipped sample parquet (bow_text / fos / 192-d
    embeddings, main.py:148-152 and client.py:344-356 of the reference): one server, two clients with different
    fields of study, CombinedTM with contextual_size taken from the data."""
    src = os.path.join(ROOT, "static", "datasets", "preprocessing_example", "s2cs_tiny_preproc.parquet")
    assert os.path.exists(src), "run aux_scripts/make_sample_data.py"
    base = _free_base()
    cfg = configparser.ConfigParser()
    cfg.read(os.path.join(ROOT, "config", "dft_params.cf"))
    cfg.set("addresses", "local", f"localhost:{base}")
    cfg.set("addresses", "base_port", str(base))
    cfg.set("federation", "server_port", str(base))
    cfg.set("federation", "time_termination", "150")
    cfg.set("federation", "client_sleep_time", "150")
    cfg.set("ntms", "n_components", "5")
    cfg.set("ntms", "hidden_sizes", "(8, 8)")
    cfg.set("ntms", "num_epochs", "1")
    cfg.set("ntms", "batch_size", "32")
    cfg.set("ntms", "num_samples", "2")
    cfg.set("ntms", "contextual_size", "192")
    cfg.set("b200", "addressing", "local")
    cfg.set("b200", "backend", "eager")
    for k in ("save_client", "save_server", "logs_client", "logs_server"):
        cfg.set("save_dir", k, str(tmp_path / "out" / k.replace("save_", "models_")))
    cfg_path = str(tmp_path / "params.cf")
    with open(cfg_path, "w") as f:
        cfg.write(f)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", PYTHONPATH=ROOT)
    common = [sys.executable, os.path.join(ROOT, "main.py"), "--config", cfg_path, "--source", src, "--local"]
    server = subprocess.Popen(common + ["--id", "0", "--min_clients_federation", "2", "--max_iters", "5",
                                        "--model_type", "ctm"], env=env, stdout=subprocess.PIPE,
                              stderr=subprocess.STDOUT, text=True)
    time.sleep(3.0)
    clients = [subprocess.Popen(common + ["--id", str(i), "--data_type", "real", "--fos", fos], env=env,
                                stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
               for i, fos in ((1, "computer_science"), (2, "medicine"))]
    try:
        out_s, _ = server.communicate(timeout=200)
        outs = [c.communicate(timeout=60)[0] for c in clients]
    finally:
        for p in [server] + clients:
            if p.poll() is None:
                p.kill()
    assert server.returncode == 0, out_s[-3000:]
    for c, o in zip(clients, outs):
        assert c.returncode == 0, o[-3000:]
    base_out = tmp_path / "out"
    assert list((base_out / "models_server").glob("global_model_*.npz"))
    for cid, n_docs in ((1, 58), (2, 64)):
        files = list(base_out.glob(f"models_client{cid}/model_{cid}_*.npz"))
        assert files, [str(p) for p in base_out.rglob("*")]
        z = np.load(str(files[0]), allow_pickle=True)
        assert z["betas"].shape[0] == 5 and z["thetas"].shape == (n_docs, 5)
