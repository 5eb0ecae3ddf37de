"""Tensor / model / optimiser (de)serialisation, result writer and config reader.

Ref: src/utils/auxiliary_functions.py — same public function names.  Differences by design:
payloads stay in the tensor's own dtype (the reference's down leg is float64), bf16/fp16/int32 are
accepted, and state-dict entries without a dedicated ``ModelUpdate`` field (third hidden layer,
label head, ...) travel in ``ModelUpdate.extra_tensors`` instead of raising.
"""
from __future__ import annotations

import configparser
import pickle
from typing import Dict, Iterable

import numpy as np
import torch
from scipy import sparse

from ..protos import federated_pb2

CHUNK_SIZE = 1024 * 1024  # 1 MB

_NP_DTYPES = {"float32": np.float32, "float64": np.float64, "int64": np.int64, "int32": np.int32,
              "float16": np.float16, "bool": np.bool_, "uint8": np.uint8}


def get_type_from_string(str_dtype: str):
    """numpy dtype for a wire dtype string (``bfloat16`` is carried as raw uint16)."""
    if str_dtype == "bfloat16":
        return np.uint16
    if str_dtype not in _NP_DTYPES:
        raise ValueError(f"dtype not defined: {str_dtype}")
    return _NP_DTYPES[str_dtype]


# ------------------------------------------------------------------------------------------------
# files
# ------------------------------------------------------------------------------------------------
def get_file_chunks(filename: str):
    with open(filename, "rb") as f:
        while True:
            piece = f.read(CHUNK_SIZE)
            if len(piece) == 0:
                return
            yield federated_pb2.Chunk(buffer=piece)


def save_chunks_to_file(chunks: Iterable, filename: str) -> None:
    with open(filename, "wb") as f:
        for chunk in chunks:
            f.write(chunk.buffer)


def unpickler(file: str):
    with open(file, "rb") as f:
        return pickle.load(f)


def pickler(file: str, ob) -> None:
    with open(file, "wb") as f:
        pickle.dump(ob, f)


def save_model_as_npz(npzfile: str, federated_model) -> None:
    """``betas``, ``thetas`` (dense, or the CSR quadruple), ``ntopics``, ``topics``.
    Ref: auxiliary_functions.py:66-99 — identical keys."""
    topics = federated_model.topics
    topics_arr = np.array(topics) if topics is not None else np.array(None, dtype=object)
    if isinstance(federated_model.thetas, sparse.csr_matrix):
        np.savez(npzfile, betas=federated_model.betas, thetas_data=federated_model.thetas.data,
                 thetas_indices=federated_model.thetas.indices, thetas_indptr=federated_model.thetas.indptr,
                 thetas_shape=federated_model.thetas.shape, ntopics=federated_model.n_components,
                 topics=topics_arr)
    else:
        np.savez(npzfile, betas=federated_model.betas, thetas=federated_model.thetas,
                 ntopics=federated_model.n_components, topics=topics_arr)


# ------------------------------------------------------------------------------------------------
# tensors
# ------------------------------------------------------------------------------------------------
def serializeTensor(tensor) -> "federated_pb2.Tensor":
    """torch tensor / numpy array -> ``Tensor`` message (raw little-endian bytes, dtype string, dims)."""
    if isinstance(tensor, torch.Tensor):
        t = tensor.detach()
        if t.dtype == torch.bfloat16:
            aux = t.cpu().contiguous().view(torch.int16).numpy().view(np.uint16)
            dtype_str = "bfloat16"
        else:
            aux = t.cpu().contiguous().numpy()
            dtype_str = str(aux.dtype)
        shape = tuple(tensor.shape)
    else:
        aux = np.ascontiguousarray(tensor)
        dtype_str = str(aux.dtype)
        shape = aux.shape
    msg = federated_pb2.Tensor(dtype=dtype_str, tensor_content=aux.tobytes())
    for i, s in enumerate(shape):
        msg.tensor_shape.dim.add(size=int(s), name="dim" + str(i))
    return msg


def deserializeNumpy(protoTensor) -> np.ndarray:
    dims = tuple(int(d.size) for d in protoTensor.tensor_shape.dim)
    arr = np.frombuffer(protoTensor.tensor_content, dtype=get_type_from_string(protoTensor.dtype))
    arr = arr.reshape(dims)
    if protoTensor.dtype == "bfloat16":
        arr = (arr.astype(np.uint32) << 16).view(np.float32)
    return arr


def deserializeTensor(protoTensor) -> torch.Tensor:
    """``Tensor`` message -> torch tensor.  Floating payloads come back as float32 (like the reference,
    whose ``torch.Tensor(np)`` always yields float32); integer payloads keep an integer dtype."""
    arr = deserializeNumpy(protoTensor)
    if np.issubdtype(arr.dtype, np.floating):
        return torch.from_numpy(np.array(arr, dtype=np.float32))
    return torch.from_numpy(np.array(arr))


# ------------------------------------------------------------------------------------------------
# model state dict <-> ModelUpdate
# ------------------------------------------------------------------------------------------------
_SPECIAL_KEYS = {"inf_net.hiddens.l_0.0.weight": "inf_net_hiddens_l00_weight",
                 "inf_net.hiddens.l_0.0.bias": "inf_net_hiddens_l_00_bias"}
_FIELDS = set(federated_pb2.MODEL_UPDATE_TENSOR_FIELDS)
_FIELD_TO_KEY = {
    "inf_net_f_mu_batchnorm_num_batches_tracked": "inf_net.f_mu_batchnorm.num_batches_tracked",
    "inf_net_f_sigma_batchnorm_num_batches_tracked": "inf_net.f_sigma_batchnorm.num_batches_tracked",
    "beta_batchnorm_num_batches_tracked": "beta_batchnorm.num_batches_tracked",
    "prior_mean": "prior_mean", "prior_variance": "prior_variance", "beta": "beta",
    "inf_net_input_layer_weight": "inf_net.input_layer.weight",
    "inf_net_input_layer_bias": "inf_net.input_layer.bias",
    "inf_net_hiddens_l00_weight": "inf_net.hiddens.l_0.0.weight",
    "inf_net_hiddens_l_00_bias": "inf_net.hiddens.l_0.0.bias",
    "inf_net_f_mu_weight": "inf_net.f_mu.weight", "inf_net_f_mu_bias": "inf_net.f_mu.bias",
    "inf_net_f_mu_batchnorm_running_mean": "inf_net.f_mu_batchnorm.running_mean",
    "inf_net_f_mu_batchnorm_running_var": "inf_net.f_mu_batchnorm.running_var",
    "inf_net_f_sigma_weight": "inf_net.f_sigma.weight", "inf_net_f_sigma_bias": "inf_net.f_sigma.bias",
    "inf_net_f_sigma_batchnorm_running_mean": "inf_net.f_sigma_batchnorm.running_mean",
    "inf_net_f_sigma_batchnorm_running_var": "inf_net.f_sigma_batchnorm.running_var",
    "beta_batchnorm_running_mean": "beta_batchnorm.running_mean",
    "beta_batchnorm_running_var": "beta_batchnorm.running_var",
    "topic_word_matrix": "topic_word_matrix",
    "inf_net_adapt_bert_weight": "inf_net.adapt_bert.weight",
    "inf_net_adapt_bert_bias": "inf_net.adapt_bert.bias",
}


def modelStateDict_to_proto(modelStateDict: dict, current_epoch: int, model_type=None):
    """State dict -> ``ModelUpdate``.  Keys with a dedicated field use it (``.`` -> ``_`` plus the two
    historical spellings); everything else goes to ``extra_tensors`` under its state-dict key."""
    mu = federated_pb2.ModelUpdate(current_epoch=int(current_epoch))
    for key, value in modelStateDict.items():
        field = _SPECIAL_KEYS.get(key, key.replace(".", "_"))
        msg = serializeTensor(value)
        if field in _FIELDS:
            getattr(mu, field).CopyFrom(msg)
        else:
            mu.extra_tensors.add(tensor_name=key, tensor=msg)
    return mu


def proto_to_modelStateDict(modelUpdate) -> Dict[str, torch.Tensor]:
    out: Dict[str, torch.Tensor] = {}
    for field, key in _FIELD_TO_KEY.items():
        if modelUpdate.HasField(field):
            t = deserializeTensor(getattr(modelUpdate, field))
            if key.endswith("num_batches_tracked"):
                t = t.reshape(()).to(torch.long)
            out[key] = t
    for upd in modelUpdate.extra_tensors:
        out[upd.tensor_name] = deserializeTensor(upd.tensor)
    return out


# ------------------------------------------------------------------------------------------------
# optimiser state dict <-> OptUpdate   (Adam only, like the reference)
# ------------------------------------------------------------------------------------------------
def optStateDict_to_proto(optStateDict: dict):
    state_ = federated_pb2.AdamUpdate.State()
    for sid, st in optStateDict.get("state", {}).items():
        state_.contentState.add(state_id=int(sid), step=serializeTensor(torch.as_tensor(st["step"])),
                                exp_avg=serializeTensor(st["exp_avg"]),
                                exp_avg_sq=serializeTensor(st["exp_avg_sq"]))
    pg = federated_pb2.AdamUpdate.ParamGroups()
    groups = optStateDict.get("param_groups", [])
    if groups:
        g = groups[0]
        pg.lr = float(g.get("lr", 0.0))
        betas = g.get("betas", (0.9, 0.999))
        pg.betas.beta1, pg.betas.beta2 = float(betas[0]), float(betas[1])
        pg.eps = float(g.get("eps", 1e-8))
        pg.weight_decay = float(g.get("weight_decay", 0.0))
        pg.amsgrad = bool(g.get("amsgrad", False))
        pg.params.extend(int(p) for p in g.get("params", []))
    return federated_pb2.OptUpdate(adamUpdate=federated_pb2.AdamUpdate(state=state_, paramGroups=pg))


def proto_to_optStateDict(optUpdate) -> dict:
    state = {}
    for cs in optUpdate.adamUpdate.state.contentState:
        state[int(cs.state_id)] = {"step": deserializeTensor(cs.step), "exp_avg": deserializeTensor(cs.exp_avg),
                                   "exp_avg_sq": deserializeTensor(cs.exp_avg_sq)}
    pg = optUpdate.adamUpdate.paramGroups
    groups = [{"lr": pg.lr, "betas": (pg.betas.beta1, pg.betas.beta2), "eps": pg.eps,
               "weight_decay": pg.weight_decay, "amsgrad": pg.amsgrad, "params": list(pg.params)}]
    return {"state": state, "param_groups": groups}


# ------------------------------------------------------------------------------------------------
# Dictionary message helpers (hyper-parameters and vocabularies)
# ------------------------------------------------------------------------------------------------
def dict_to_proto(params: dict):
    """Python dict -> ``Dictionary`` (str / None / float / bool / int / tuple-of-int values).  Values of
    other types (e.g. the ``id2token`` map) are skipped; the reference re-appends the previous pair in
    that case (server.py:263-267)."""
    d = federated_pb2.Dictionary()
    V = federated_pb2.Dictionary.Pair.Value
    for key, value in params.items():
        if isinstance(value, bool):
            v = V(bvalue=value)
        elif isinstance(value, str):
            v = V(svalue=value)
        elif value is None:
            v = V(svalue="None")
        elif isinstance(value, float):
            v = V(fvalue=value)
        elif isinstance(value, int):
            v = V(ivalue=value)
        elif isinstance(value, tuple):
            t = federated_pb2.Tuple()
            for el in value:
                t.values.add(ivalue=int(el))
            v = V(tvalue=t)
        else:
            continue
        d.pairs.add(key=key, value=v)
    return d


def proto_to_dict(d) -> dict:
    out = {}
    for pair in d.pairs:
        which = pair.value.WhichOneof("oneof_values")
        if which == "svalue":
            out[pair.key] = None if pair.value.svalue == "None" else pair.value.svalue
        elif which == "ivalue":
            out[pair.key] = pair.value.ivalue
        elif which == "fvalue":
            out[pair.key] = pair.value.fvalue
        elif which == "tvalue":
            out[pair.key] = tuple(el.ivalue for el in pair.value.tvalue.values)
        elif which == "bvalue":
            out[pair.key] = pair.value.bvalue
    return out


def vocab_to_proto(vocab: Dict[str, int]):
    d = federated_pb2.Dictionary()
    V = federated_pb2.Dictionary.Pair.Value
    P = federated_pb2.Dictionary.Pair
    d.pairs.extend(P(key=k, value=V(ivalue=int(i))) for k, i in vocab.items())
    return d


def proto_to_vocab(d) -> Dict[str, int]:
    return {pair.key: pair.value.ivalue for pair in d.pairs}


# ------------------------------------------------------------------------------------------------
# config
# ------------------------------------------------------------------------------------------------
_INT_KEYS = {"n_components", "num_iterations", "batch_size", "num_threads", "optimize_interval", "num_epochs",
             "num_samples", "num_data_loader_workers", "contextual_size", "max_features"}
_FLOAT_KEYS = {"thetas_thr", "doc_topic_thr", "alpha", "dropout", "lr", "momentum", "topic_prior_mean",
               "dropout_in", "dropout_out"}
_BOOL_KEYS = {"learn_priors", "reduce_on_plateau", "verbose"}


def read_config_experiments(file_path: str, skip=()) -> dict:
    """Flattens every section of the INI file into one dict with per-key type coercion.
    Ref: auxiliary_functions.py:387-438 (same coercion table; ``max_features`` and the two
    ``dropout_in/out`` keys are additionally typed)."""
    config = configparser.ConfigParser()
    config.read(file_path)
    out = {}
    for section in config.sections():
        if section in skip:
            continue
        for option in config.options(section):
            value = config.get(section, option)
            if option in _INT_KEYS:
                out[option] = int(value)
            elif option in _FLOAT_KEYS:
                out[option] = float(value)
            elif option == "labels":
                out[option] = ""
            elif option == "topic_prior_variance":
                out[option] = None
            elif option in _BOOL_KEYS:
                out[option] = value == "True"
            elif option == "hidden_sizes":
                out[option] = tuple(int(x) for x in value.strip()[1:-1].split(",") if x.strip())
            else:
                out[option] = value
    return out


def convert_topic_word_to_init_size(vocab_size: int, model, model_type: str, ntopics: int, id2token,
                                    all_words) -> np.ndarray:
    """Topic-word distribution re-indexed to the full synthetic vocabulary (zeros for words the model
    never saw), rows renormalised.  Ref: auxiliary_functions.py:441-483."""
    if model_type in ("avitm", "ctm"):
        w_t_distrib = np.zeros((ntopics, vocab_size), dtype=np.float64)
        wd = model.get_topic_word_distribution()
        pos = {w: i for i, w in enumerate(all_words)}
        items = id2token.items() if isinstance(id2token, dict) else enumerate(id2token)
        for j, tok in items:
            if tok in pos:
                w_t_distrib[:, pos[tok]] = wd[:, int(j)]
        s = w_t_distrib.sum(axis=1, keepdims=True)
        s[s == 0] = 1.0
        return w_t_distrib / s
    raise ValueError("method not specified")
