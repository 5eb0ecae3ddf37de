"""Record of one client of the federation as seen by the server.
Ref: src/federation/federation_client.py:10-125 (same attribute names and static helpers)."""
from __future__ import annotations

from typing import Dict, List, Optional


class FederationClient:
    def __init__(self, federation_key: str, tensors: Optional[Dict] = None, current_epoch: int = -1,
                 current_id_msg: int = -2, num_max_epochs: int = -3, vocab_sent: bool = False,
                 ready_for_training: bool = False, client_id: int = -1, nr_samples: int = 0) -> None:
        self.federation_key = federation_key      # gRPC peer string of the consensus-stage connection
        self.tensors = tensors if tensors is not None else {}
        self.current_mb = -1
        self.current_epoch = current_epoch
        self.current_id_msg = current_id_msg
        self.num_max_epochs = num_max_epochs
        self.vocab_sent = vocab_sent
        self.ready_for_training = ready_for_training
        self.client_id = client_id
        self.nr_samples = nr_samples
        self.num_max_iter = -3                    # reference default (federation_client.py:24)
        self.finished = False                     # reached num_epochs (server stops polling it)
        self.last_seen = 0.0                      # heartbeat: time of the last successful RPC
        self.failures = 0

    def set_num_max_iter(self, num_max_iter: int) -> None:
        """Maximum number of training iterations announced by the client (federation_client.py:32-46)."""
        self.num_max_iter = num_max_iter

    def update_client_state(self, tensors: Dict, current_mb: int, current_epoch: int, current_id_msg) -> None:
        self.tensors = tensors
        self.current_mb = current_mb
        self.current_epoch = current_epoch        # (the reference stores a 1-tuple here by accident)
        self.current_id_msg = current_id_msg

    # -- helpers over a list of clients -------------------------------------------------------
    @staticmethod
    def get_pos_by_key(key: str, federation_clients: List["FederationClient"]) -> int:
        for pos, c in enumerate(federation_clients):
            if c.federation_key == key:
                return pos
        return -1

    @staticmethod
    def get_pos_by_id(client_id: int, federation_clients: List["FederationClient"]) -> int:
        for pos, c in enumerate(federation_clients):
            if c.client_id == client_id:
                return pos
        return -1

    @staticmethod
    def set_can_start_training(key: str, federation_clients: List["FederationClient"], value: bool = True) -> None:
        pos = FederationClient.get_pos_by_key(key, federation_clients)
        if pos >= 0:
            federation_clients[pos].ready_for_training = value

    @staticmethod
    def set_id_by_key(key: str, federation_clients: List["FederationClient"], id: int) -> None:
        pos = FederationClient.get_pos_by_key(key, federation_clients)
        if pos >= 0:
            federation_clients[pos].client_id = id
