"""Federation client: stage-1 driver (``Client``) and the client-side gRPC server the federation
server pulls from during training (``FederatedClientServer``).

Ref: src/federation/client.py:43-532 — same class names, constructor arguments and RPC behaviour.
Differences: the local corpus is vectorised against the consensus vocabulary into a *sparse* CSR matrix
(the reference calls ``.toarray()``, client.py:461-462); ``grads_to_share`` entries a model does not have
are ignored; the client wakes up when the server's STOP arrives instead of sleeping a fixed week.
"""
from __future__ import annotations

import os
import threading
import time
from concurrent import futures
from typing import Optional, Union

import grpc
import numpy as np
from sklearn.feature_extraction.text import CountVectorizer

from ..federated.federated_avitm import FederatedAVITM
from ..federated.federated_ctm import FederatedCTM
from ..models.datasets import BOWDataset
from ..models.data_preparation import TopicModelDataPreparation
from ..protos import federated_pb2, federated_pb2_grpc
from ..utils.auxiliary_functions import (proto_to_dict, proto_to_modelStateDict, proto_to_optStateDict,
                                         proto_to_vocab, serializeTensor, vocab_to_proto)
from .server import make_logger


class FederatedClientServer(federated_pb2_grpc.FederationServerServicer):
    """Serves ``getGradient`` / ``sendAggregatedTensor`` for one client."""

    def __init__(self, local_model: Union[FederatedAVITM, FederatedCTM], train_data, id: int,
                 save_client: str, logger=None, checkpoint_path: str = None, checkpoint_every: int = 0,
                 resume: bool = False) -> None:
        self._local_model = local_model
        self.id = id
        self._save_client = save_client
        self.global_epoch = -1
        self.stopped = threading.Event()
        self._lock = threading.Lock()
        self._logger = logger if logger else make_logger("ClientServer", None)
        self._local_model.preFit(train_data)
        # mid-training checkpoints (utils/checkpoint.py): every `checkpoint_every` aggregated updates; `resume`
        # restores model, optimiser, bookkeeping and RNG state before the first pull
        self._ckpt_path, self._ckpt_every, self._n_updates = checkpoint_path, int(checkpoint_every or 0), 0
        if resume and checkpoint_path and os.path.exists(checkpoint_path):
            from ..utils.checkpoint import load_training_checkpoint

            load_training_checkpoint(checkpoint_path, self._local_model, getattr(self._local_model, "train_loader", None))
            self._logger.info("-- -- Client %s resumed from %s", str(self.id), checkpoint_path)
        self._last_iter, self._last_reply = None, None
        self._logger.info("FederatedClientServer object initialized")

    def getGradient(self, request, context):
        with self._lock:
            # Idempotent per global iteration: a server-side retry of a timed-out pull must not train a second
            # time on the same minibatch (the optimiser would advance twice while the counters advance once).
            if self._last_reply is not None and request.iter == self._last_iter:
                return self._last_reply
            self.global_epoch = request.iter
            gr_upt = self._local_model.train_mb_delta()
            header = federated_pb2.MessageHeader(
                id_request="ID" + str(self.id) + "_" + str(round(time.time())),
                message_type=federated_pb2.MessageType.CLIENT_TENSOR_SEND)
            metadata = federated_pb2.MessageAdditionalData(
                current_mb=int(gr_upt["current_mb"]), current_epoch=int(gr_upt["current_epoch"]),
                num_max_epochs=int(gr_upt["num_epochs"]), id_machine=int(self.id))
            updates = [federated_pb2.Update(tensor_name=k, tensor=serializeTensor(v))
                       for k, v in gr_upt.items() if k not in ("current_mb", "current_epoch", "num_epochs")]
            self._logger.info("-- -- Client %s sent gradient update for global epoch %s", str(self.id),
                              self.global_epoch)
            reply = federated_pb2.ClientTensorRequest(header=header, metadata=metadata, updates=updates)
            self._last_iter, self._last_reply = request.iter, reply
            return reply

    def sendAggregatedTensor(self, request, context):
        mt = request.header.message_type
        if mt == federated_pb2.MessageType.SERVER_AGGREGATED_TENSOR_SEND:
            with self._lock:
                self._logger.info("-- -- Client-server %s received aggregated tensor for global epoch %s",
                                  str(self.id), self.global_epoch)
                state = proto_to_modelStateDict(request.nndata.modelUpdate)
                self._local_model.deltaUpdateFit(modelStateDict=state, save_dir=self._save_client)
                self._n_updates += 1
                if self._ckpt_path and self._ckpt_every > 0 and self._n_updates % self._ckpt_every == 0 \
                        and not self._local_model.finished:
                    from ..utils.checkpoint import save_training_checkpoint

                    save_training_checkpoint(self._ckpt_path, self._local_model,
                                             getattr(self._local_model, "train_loader", None),
                                             extra={"updates": self._n_updates, "global_epoch": int(self.global_epoch)})
            # the ack tells the server whether this client has just reached num_epochs (it rolls its epoch counter
            # and saves its results while installing the aggregate), so that it is not pulled again
            done = ";finished" if self._local_model.finished else ""
            header = federated_pb2.MessageHeader(id_request=str(self.global_epoch) + done,
                                                 message_type=federated_pb2.MessageType.CLIENT_CONFIRM_RECEIVED)
            return federated_pb2.ClientReceivedResponse(header=header)
        if mt == federated_pb2.MessageType.SERVER_STOP_TRAINING_REQUEST:
            self._logger.info("-- -- Client-server %s received request for finished training", str(self.id))
            with self._lock:
                if not self._local_model.finished:
                    self._logger.info("-- -- Getting results at client side...")
                    self._local_model.get_results_model(self._save_client)
            self.stopped.set()
            header = federated_pb2.MessageHeader(message_type=federated_pb2.MessageType.CLIENT_CONFIRM_RECEIVED)
            return federated_pb2.ClientReceivedResponse(header=header)
        context.abort(grpc.StatusCode.INVALID_ARGUMENT, f"unexpected message type {mt}")


class Client:
    """Runs the whole client life cycle from its constructor, like the reference."""

    def __init__(self, id: int, stub, local_corpus, data_type: str, opts_server: list, save_client: str,
                 logs_client: str, sleep_time: int, base_port: int,
                 grads_to_share=["prior_mean", "prior_variance", "beta"], logger=None, share: str = "param",
                 backend: str = "auto", device=None, checkpoint_path: str = None, checkpoint_every: int = 0,
                 resume: bool = False) -> None:
        self.id = id
        self._ckpt = (checkpoint_path, checkpoint_every, resume)
        self._stub = stub
        self._save_client = save_client
        self._sleep_time = sleep_time
        self._base_port = base_port
        self._grads_to_share = grads_to_share
        self._share = share
        self._backend = backend
        self._device = device
        self._logger = logger if logger else make_logger(f"Client{id}", logs_client)

        self._local_corpus, self._local_embeddings = self.__get_local_corpus(data_type, local_corpus)
        self._local_model_type = None
        self._model_parameters = None
        self.local_model = None
        self.train_data = None
        self._global_epoch = -3
        self.client_server = None
        self.federated_server: Optional[FederatedClientServer] = None

        self.__send_local_vocab()
        self.__wait_for_agreed_vocab_NN()
        self.__start_client_server(opts_server, save_client)
        self.__send_ready_for_training()

    # ------------------------------------------------------------------------------------------
    def __get_local_corpus(self, data_type: str, corpus):
        """Token lists (synthetic) or a DataFrame with ``bow_text`` [+ ``embeddings``] (real)."""
        embeddings = None
        if data_type == "synthetic":
            docs = [" ".join(corpus[i]) for i in np.arange(len(corpus))]
        else:
            col = corpus["bow_text"].values.tolist()
            docs = [d if isinstance(d, str) else " ".join(d) for d in col]
            if "embeddings" in list(corpus.columns.values):
                embeddings = corpus["embeddings"].values
                if len(embeddings) and isinstance(embeddings[0], str):
                    embeddings = np.array([np.array(el.split(), dtype=np.float32) for el in embeddings])
                else:
                    embeddings = np.stack([np.asarray(e, dtype=np.float32) for e in embeddings])
        return docs, embeddings

    def __prepare_vocab_to_send(self) -> dict:
        cv = CountVectorizer(input="content", lowercase=True, stop_words="english", binary=False)
        cv.fit(self._local_corpus)
        return {k: int(v) for k, v in cv.vocabulary_.items()}

    def __send_local_vocab(self) -> None:
        vocab_dict = self.__prepare_vocab_to_send()
        req = federated_pb2.DictRequest(vocab=vocab_to_proto(vocab_dict), client_id=self.id,
                                        nr_samples=len(self._local_corpus))
        if self._stub:
            response = self._stub.sendLocalDic(req)
            self._logger.info("-- -- Client %s vocab is being sent to server.", str(self.id))
            if response.length == len(vocab_dict):
                self._logger.info("-- -- Server received correctly vocab from client %s.", str(self.id))

    def __wait_for_agreed_vocab_NN(self) -> None:
        self._logger.info("-- -- Client %s receiving consensus vocab and initial NN.", str(self.id))
        response = self._stub.sendGlobalDicAndInitialNN(federated_pb2.Empty())
        self._model_parameters = proto_to_dict(response.model_params)
        self._local_model_type = response.model_type
        vocab = proto_to_vocab(response.dic[0])
        self._global_vocab = CountVectorizer(vocabulary=vocab)
        # sparse document-term matrix over the consensus vocabulary (never densified)
        train_bow = self._global_vocab.transform(self._local_corpus)
        idx2token = self._global_vocab.get_feature_names_out()
        self._model_parameters["input_size"] = len(idx2token)
        id2token = dict(enumerate(idx2token))
        self._model_parameters["id2token"] = id2token
        self._model_parameters.setdefault("seed", 1000 + int(self.id))
        if self._local_model_type == "avitm":
            self.train_data = BOWDataset(train_bow, idx2token)
            self.local_model = FederatedAVITM(self._model_parameters, self._grads_to_share, self._logger,
                                              share=self._share, backend=self._backend, device=self._device)
        elif self._local_model_type == "ctm":
            if self._local_embeddings is None:
                raise ValueError("model_type 'ctm' needs an `embeddings` column in the client's corpus")
            qt = TopicModelDataPreparation()
            qt.vectorizer, qt.id2token, qt.vocab = self._global_vocab, id2token, idx2token
            self._model_parameters["contextual_size"] = int(self._local_embeddings.shape[1])
            self.train_data = qt.load(contextualized_embeddings=self._local_embeddings,
                                      bow_embeddings=train_bow, id2token=id2token)
            self.local_model = FederatedCTM(self._model_parameters, self._grads_to_share, self._logger,
                                            share=self._share, backend=self._backend, device=self._device)
        else:
            raise ValueError("Provided underlying model not supported")
        state = proto_to_modelStateDict(response.initialNN.modelUpdate)
        own = self.local_model.model.state_dict()
        compatible = {k: v for k, v in state.items() if k in own and tuple(own[k].shape) == tuple(v.shape)}
        self.local_model.model.load_state_dict(compatible, strict=False)
        self.local_model.parameters_changed()
        opt_state = proto_to_optStateDict(response.initialNN.optUpdate)
        if opt_state["state"]:
            self.local_model.optimizer.load_state_dict(opt_state)
        self._logger.info("-- -- Client %s initialized local model appropiately.", str(self.id))

    def __start_client_server(self, opts_server: list, save_client: str) -> None:
        self.client_server = grpc.server(futures.ThreadPoolExecutor(max_workers=4), options=opts_server)
        self.federated_server = FederatedClientServer(self.local_model, self.train_data, self.id, save_client,
                                                      self._logger, checkpoint_path=self._ckpt[0],
                                                      checkpoint_every=self._ckpt[1], resume=self._ckpt[2])
        federated_pb2_grpc.add_FederationServerServicer_to_server(self.federated_server, self.client_server)
        self.client_server.add_insecure_port("[::]:" + str(self._base_port + self.id))
        self.client_server.start()
        self._logger.info(f"-- -- Starting Client-server at {str(self._base_port + self.id)}...")

    def __send_ready_for_training(self) -> None:
        self._logger.info("-- -- Client %s sending ready for training phase.", str(self.id))
        header = federated_pb2.MessageHeader(id_request="ID" + str(self.id) + "_" + str(round(time.time())),
                                             message_type=federated_pb2.MessageType.CLIENT_READY_FOR_TRAINING)
        request = federated_pb2.ClientTensorRequest(header=header)
        if self._stub:
            self._stub.trainFederatedModel(request)
            self._logger.info(f"-- -- Client {self.id} stop behaving as client. Waiting for training to finish...")
            self.federated_server.stopped.wait(timeout=self._sleep_time)
            self.client_server.stop(grace=1.0)
