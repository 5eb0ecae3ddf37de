#!/usr/bin/env python
"""Command line entry point: federation server (``--id 0``), gRPC client (``--id N``) or an in-box
NVLink federation with one client per GPU (``--fabric N``).

Ref: main.py:27-295 of the reference — same flags and defaults, same config file
(``config/dft_params.cf``), same output layout (``static/output_models/{server,client<id>}``,
``static/logs/...``, ``YYYYMMDD`` suffixes).  ``WORKDIR`` comes from the environment variable of that
name (the reference's Docker images export it) and falls back to this file's directory instead of the
hard-coded ``/workspace``.
"""
from __future__ import annotations

import argparse
import configparser
import datetime as DT
import os
import pathlib
import sys
from concurrent import futures

import numpy as np

WORKDIR = os.environ.get("WORKDIR", os.path.dirname(os.path.abspath(__file__)))
if WORKDIR not in sys.path:
    sys.path.insert(0, WORKDIR)


def _abs(path: str) -> str:
    return path if os.path.isabs(path) else os.path.join(WORKDIR, path)


def start_server(min_num_clients, model_type, max_iters, opts_server, opts_client, address, training_params,
                 save_server, logs_server, grads_to_share, server_port=50051, time_termination=1200,
                 base_port=50051, agg="param", addressing="docker"):
    """Starts the gRPC federation server and blocks until training finished (or ``time_termination``)."""
    import grpc

    from gfedntm_b200.federation.server import FederatedServer
    from gfedntm_b200.protos import federated_pb2_grpc

    client_server_addres = "gfedntm-client" if (addressing == "docker" and address.startswith("gfedntm-server")) \
        else "localhost"
    os.makedirs(save_server, exist_ok=True)
    os.makedirs(logs_server, exist_ok=True)
    save_server = os.path.join(save_server, f"global_model_{DT.datetime.now().strftime('%Y%m%d')}")
    logs_server = os.path.join(logs_server, f"logs_{DT.datetime.now().strftime('%Y%m%d')}.txt")
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=32), options=opts_server)
    federated_server = FederatedServer(min_num_clients=min_num_clients, model_params={**training_params},
                                       model_type=model_type, max_iters=max_iters, opts_client=opts_client,
                                       save_server=save_server, logs_server=logs_server,
                                       grads_to_share=grads_to_share, client_server_addres=client_server_addres,
                                       base_port=base_port, agg=agg)
    federated_pb2_grpc.add_FederationServicer_to_server(federated_server, server)
    server.add_insecure_port("[::]:" + str(server_port))
    server.start()
    federated_server.training_finished.wait(timeout=time_termination)
    server.stop(grace=2.0)
    if federated_server.training_error is not None:
        raise federated_server.training_error
    return federated_server


def load_client_corpus(data_type: str, source: str, id_client: int, fos: str):
    """Synthetic: ``documents[id-1]`` of the npz; real: parquet rows whose ``fos`` column matches."""
    if data_type == "synthetic":
        data = np.load(source, allow_pickle=True)
        return data["documents"][id_client - 1]
    if data_type == "real":
        import pandas as pd

        df = pd.read_parquet(pathlib.Path(source))
        if "fos" in df.columns:
            df = df[df["fos"] == fos]
        return df
    raise ValueError("Specified data type not supported")


def start_client(id_client, data_type, fos, source, address, opts_client, opts_server, save_client, logs_client,
                 client_sleep_time, grads_to_share, base_port=50051, agg="param", backend="auto", checkpoint_dir=None,
                 checkpoint_every=0, resume=False):
    import grpc

    from gfedntm_b200.federation.client import Client
    from gfedntm_b200.protos import federated_pb2_grpc

    corpus = load_client_corpus(data_type, source, id_client, fos)
    save_dir = f"{save_client}{id_client}"
    logs_dir = f"{logs_client}{id_client}"
    os.makedirs(save_dir, exist_ok=True)
    os.makedirs(logs_dir, exist_ok=True)
    save_path = os.path.join(save_dir, f"model_{id_client}_{DT.datetime.now().strftime('%Y%m%d')}")
    logs_path = os.path.join(logs_dir, f"logs_{DT.datetime.now().strftime('%Y%m%d')}.txt")
    with grpc.insecure_channel(address, options=opts_client) as channel:
        stub = federated_pb2_grpc.FederationStub(channel)
        return Client(id=id_client, stub=stub, local_corpus=corpus, data_type=data_type, opts_server=opts_server,
                      save_client=save_path, logs_client=logs_path, sleep_time=client_sleep_time,
                      base_port=base_port, grads_to_share=grads_to_share,
                      share="grad" if agg == "grad" else "param", backend=backend,
                      checkpoint_path=(os.path.join(checkpoint_dir, f"client{id_client}.ckpt") if checkpoint_dir else None),
                      checkpoint_every=checkpoint_every, resume=resume)


def main(argv=None):
    parser = argparse.ArgumentParser()
    # client parameters
    parser.add_argument("--id", type=int, default=0, help="Client ID. If not provided, the server (ID=0) is started.")
    parser.add_argument("--source", type=str, default=os.path.join(WORKDIR, "static/datasets/synthetic.npz"),
                        help="Path to the training data.")
    parser.add_argument("--data_type", type=str, default="synthetic", help="synthetic or real")
    parser.add_argument("--fos", type=str, default="computer_science", help="Category")
    # server parameters
    parser.add_argument("--min_clients_federation", type=int, default=1,
                        help="Minimum number of client that are necessary for starting a federation. "
                             "This parameter only affects the server.")
    parser.add_argument("--model_type", type=str, default="avitm", help="Underlying model type: avitm/ctm.")
    parser.add_argument("--max_iters", type=int, default=25000,
                        help="Maximum number of global iterations to train the federated topic model.")
    # gfedntm_b200 additions
    parser.add_argument("--config", type=str, default=os.path.join(WORKDIR, "config/dft_params.cf"))
    parser.add_argument("--fabric", type=int, default=0,
                        help="Run an in-box federation with this many clients, one per GPU, aggregated by the "
                             "fused NVLink kernel (no gRPC data plane).")
    parser.add_argument("--local", action="store_true", help="Use the [addresses] local entry (no Docker DNS).")
    parser.add_argument("--checkpoint_every", type=int, default=0,
                        help="Write a mid-training checkpoint every N global steps (clients and --fabric ranks).")
    parser.add_argument("--checkpoint_dir", type=str, default=os.path.join(WORKDIR, "static/checkpoints"),
                        help="Directory of the mid-training checkpoints.")
    parser.add_argument("--resume", action="store_true", help="Continue from the checkpoint in --checkpoint_dir.")
    parser.add_argument("--max_restarts", type=int, default=0,
                        help="--fabric: restart all ranks from the last checkpoint when one of them dies, at most this "
                             "many times (needs --checkpoint_every).")
    args = parser.parse_args(argv)

    from gfedntm_b200.utils.auxiliary_functions import read_config_experiments

    training_params = read_config_experiments(args.config)
    config = configparser.ConfigParser()
    config.read(args.config)
    addressing = config.get("b200", "addressing", fallback="docker")
    if args.local:
        addressing = "local"
    address = config.get("addresses", "local" if addressing == "local" else "docker")
    base_port = int(config.get("addresses", "base_port"))
    agg = config.get("b200", "agg", fallback="param")
    backend = config.get("b200", "backend", fallback="auto")

    max_len = int(config.get("grpc", "max_message_length"))
    max_in = int(config.get("grpc", "max_inbound_message_size"))
    max_meta = int(config.get("grpc", "max_inbound_metadata_size"))
    opts_client = [("grpc.max_message_length", max_len), ("grpc.max_send_message_length", max_len),
                   ("grpc.max_receive_message_length", max_len), ("grpc.max_inbound_message_size", max_in),
                   ("grpc.max_inbound_metadata_size", max_meta), ("grpc.max_metadata_size", max_meta)]
    opts_server = [("grpc.max_send_message_length", max_len), ("grpc.max_receive_message_length", max_len),
                   ("grpc.keepalive_time_ms", int(config.get("grpc", "keepalive_time_ms"))),
                   ("grpc.keepalive_timeout_ms", int(config.get("grpc", "keepalive_timeout_ms"))),
                   ("grpc.keepalive_permit_without_calls",
                    config.get("grpc", "keepalive_permit_without_calls") == "True"),
                   ("grpc.http2.max_ping_strikes", int(config.get("grpc", "max_ping_strikes")))]
    grads_to_share = config.get("federation", "grads_to_share").split(",")
    save_client = _abs(config.get("save_dir", "save_client"))
    save_server = _abs(config.get("save_dir", "save_server"))
    logs_client = _abs(config.get("save_dir", "logs_client"))
    logs_server = _abs(config.get("save_dir", "logs_server"))

    if args.fabric > 0:
        from gfedntm_b200.federated.fabric_run import run_fabric_federation

        return run_fabric_federation(n_clients=args.fabric, source=args.source, data_type=args.data_type,
                                     fos=args.fos, model_type=args.model_type, training_params=training_params,
                                     max_iters=args.max_iters, save_client=save_client, save_server=save_server,
                                     logs_client=logs_client, agg="grad" if agg == "grad" else "param",
                                     grads_to_share=grads_to_share, checkpoint_dir=args.checkpoint_dir,
                                     checkpoint_every=args.checkpoint_every, resume=args.resume,
                                     max_restarts=args.max_restarts)
    if args.id == 0:
        print("Starting server with", args.min_clients_federation,
              "as minimum number of clients to start the federation.")
        return start_server(min_num_clients=args.min_clients_federation, model_type=args.model_type,
                            max_iters=args.max_iters, opts_server=opts_server, opts_client=opts_client,
                            address=address, training_params=training_params, save_server=save_server,
                            logs_server=logs_server, grads_to_share=grads_to_share,
                            server_port=int(config.get("federation", "server_port")),
                            time_termination=int(config.get("federation", "time_termination")),
                            base_port=base_port, agg=agg, addressing=addressing)
    print("Starting client with id ", args.id)
    return start_client(id_client=args.id, data_type=args.data_type, fos=args.fos, source=args.source,
                        address=address, opts_client=opts_client, opts_server=opts_server,
                        save_client=save_client, logs_client=logs_client,
                        client_sleep_time=int(config.get("federation", "client_sleep_time")),
                        grads_to_share=grads_to_share, base_port=base_port, agg=agg, backend=backend,
                        checkpoint_dir=args.checkpoint_dir, checkpoint_every=args.checkpoint_every, resume=args.resume)


if __name__ == "__main__":
    main()
