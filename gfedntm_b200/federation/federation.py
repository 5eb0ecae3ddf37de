"""Thread-safe registry of the clients of a federation.
Ref: src/federation/federation.py:14-170.

The reference keys its records by gRPC peer string and reference-counts connections through RPC
termination callbacks; records are in practice never removed and the list doubles as the training
roster.  The same API is kept (``connect_*`` / ``disconnect`` / ``getClients``); on top of it clients are
also addressable by their numeric id, which is what a reconnecting client keeps across channels.
"""
from __future__ import annotations

import logging
import threading
from typing import Dict, List

from .federation_client import FederationClient


class Federation:
    def __init__(self, logger=None) -> None:
        self.logger = logger if logger is not None else logging.getLogger("Federation")
        self.federation_lock = threading.RLock()
        self.federation: Dict[str, int] = {}            # peer string -> open-connection count
        self.federation_clients: List[FederationClient] = []

    # -- connections --------------------------------------------------------------------------
    def _connect(self, client: str, **fields) -> None:
        with self.federation_lock:
            if client not in self.federation:
                self.federation[client] = 1
                self.federation_clients.append(FederationClient(federation_key=client, **fields))
            else:
                self.federation[client] += 1
                pos = FederationClient.get_pos_by_key(client, self.federation_clients)
                for k, v in fields.items():
                    setattr(self.federation_clients[pos], k, v)

    def connect_consensus(self, client: str) -> None:
        """A client opened the RPC that delivers its local vocabulary."""
        self._connect(client)

    def connect_waiting_or_consensus(self, client: str, waiting: bool) -> None:
        """A client is waiting for the consensus vocabulary (``waiting=False``) or for training to start
        (``waiting=True``)."""
        self._connect(client)

    def connect_update(self, client: str, gradient: dict, current_mb: int, current_epoch: int,
                       current_id_msg, max_epochs: int) -> None:
        with self.federation_lock:
            self._connect(client)
            pos = FederationClient.get_pos_by_key(client, self.federation_clients)
            c = self.federation_clients[pos]
            c.update_client_state(gradient, current_mb, current_epoch, current_id_msg)
            c.set_num_max_iter(max_epochs)
            c.num_max_epochs = max_epochs

    def disconnect(self, client: str) -> None:
        """Called when an RPC of ``client`` terminates.  The record itself is kept: it is the roster."""
        with self.federation_lock:
            if client not in self.federation:
                return
            self.federation[client] -= 1
            if self.federation[client] <= 0:
                self.federation[client] = 0

    def merge_duplicate_ids(self) -> None:
        """Collapse records that share a client id (a client that used more than one channel during the
        consensus stage), keeping the union of their flags."""
        with self.federation_lock:
            by_id: Dict[int, FederationClient] = {}
            merged: List[FederationClient] = []
            for c in self.federation_clients:
                if c.client_id < 0 or c.client_id not in by_id:
                    if c.client_id >= 0:
                        by_id[c.client_id] = c
                    merged.append(c)
                else:
                    k = by_id[c.client_id]
                    k.vocab_sent |= c.vocab_sent
                    k.ready_for_training |= c.ready_for_training
                    k.nr_samples = max(k.nr_samples, c.nr_samples)
            self.federation_clients = merged

    def getClients(self) -> List[str]:
        with self.federation_lock:
            return list(self.federation.keys())

    def client_by_id(self, client_id: int) -> FederationClient:
        with self.federation_lock:
            pos = FederationClient.get_pos_by_id(client_id, self.federation_clients)
            if pos < 0:
                raise KeyError(client_id)
            return self.federation_clients[pos]
