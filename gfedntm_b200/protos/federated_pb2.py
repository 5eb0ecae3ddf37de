"""Message classes of package ``federated`` (see ``federated.proto`` next to this file).

``grpc_tools`` / ``protoc`` are not required: the file descriptor is assembled here with
``descriptor_pb2`` and the classes are created through the protobuf runtime.  Field numbers and
names match the reference's generated ``federated_pb2`` so the two are wire compatible.
"""
from __future__ import annotations

from google.protobuf import descriptor_pb2 as _d
from google.protobuf import descriptor_pool as _pool
from google.protobuf import message_factory as _mf

_F = _d.FieldDescriptorProto
_T = {"string": _F.TYPE_STRING, "bytes": _F.TYPE_BYTES, "int32": _F.TYPE_INT32, "int64": _F.TYPE_INT64,
      "uint32": _F.TYPE_UINT32, "float": _F.TYPE_FLOAT, "bool": _F.TYPE_BOOL}

PACKAGE = "federated"


def _field(msg, name, number, ftype, *, repeated=False, oneof=None, optional=False):
    f = msg.field.add()
    f.name, f.number = name, number
    f.label = _F.LABEL_REPEATED if repeated else _F.LABEL_OPTIONAL
    if ftype in _T:
        f.type = _T[ftype]
    elif ftype.startswith("enum:"):
        f.type = _F.TYPE_ENUM
        f.type_name = "." + PACKAGE + "." + ftype[5:]
    else:
        f.type = _F.TYPE_MESSAGE
        f.type_name = "." + PACKAGE + "." + ftype
    if oneof is not None:
        f.oneof_index = oneof
    if optional:
        msg.oneof_decl.add().name = "_" + name
        f.oneof_index = len(msg.oneof_decl) - 1
        f.proto3_optional = True
    return f


def _build_file() -> _d.FileDescriptorProto:
    fd = _d.FileDescriptorProto()
    fd.name = "gfedntm_b200/federated.proto"
    fd.package = PACKAGE
    fd.syntax = "proto3"

    def msg(name, parent=None):
        m = (parent.nested_type if parent is not None else fd.message_type).add()
        m.name = name
        return m

    msg("Empty")

    ts = msg("TensorShape")
    dim = msg("Dim", ts)
    _field(dim, "size", 1, "int64")
    _field(dim, "name", 2, "string", optional=True)
    _field(ts, "dim", 2, "TensorShape.Dim", repeated=True)

    t = msg("Tensor")
    _field(t, "tensor_shape", 1, "TensorShape")
    _field(t, "dtype", 2, "string")
    _field(t, "tensor_content", 3, "bytes")

    u = msg("Update")
    _field(u, "tensor_name", 1, "string")
    _field(u, "tensor", 2, "Tensor")

    md = msg("MessageAdditionalData")
    for i, n in enumerate(("current_mb", "current_epoch", "num_max_epochs", "id_machine"), 1):
        _field(md, n, i, "uint32")

    en = fd.enum_type.add()
    en.name = "MessageType"
    for i, n in enumerate(("CLIENT_TENSOR_SEND", "CLIENT_CONFIRM_RECEIVED", "CLIENT_READY_FOR_TRAINING",
                           "SERVER_AGGREGATED_TENSOR_SEND", "SERVER_CONFIRM_RECEIVED",
                           "SERVER_STOP_TRAINING_REQUEST")):
        v = en.value.add()
        v.name, v.number = n, i

    mh = msg("MessageHeader")
    _field(mh, "id_request", 1, "string", optional=True)
    _field(mh, "id_response", 2, "string", optional=True)
    _field(mh, "id_to_request", 3, "string", optional=True)
    _field(mh, "message_type", 4, "enum:MessageType")

    ctr = msg("ClientTensorRequest")
    _field(ctr, "header", 1, "MessageHeader")
    _field(ctr, "metadata", 2, "MessageAdditionalData")
    _field(ctr, "updates", 3, "Update", repeated=True)

    sat = msg("ServerAggregatedTensorRequest")
    _field(sat, "header", 1, "MessageHeader")
    _field(sat, "metadata", 2, "MessageAdditionalData")
    sat.oneof_decl.add().name = "oneof_values"
    _field(sat, "data", 3, "Update", oneof=0)
    _field(sat, "nndata", 4, "NNUpdate", oneof=0)

    for n in ("ClientReceivedResponse", "ServerReceivedResponse"):
        r = msg(n)
        _field(r, "header", 1, "MessageHeader")
        _field(r, "metadata", 2, "MessageAdditionalData")

    _field(msg("Chunk"), "buffer", 1, "bytes")
    _field(msg("Request"), "name", 1, "string")
    _field(msg("Reply"), "length", 1, "int32")

    tp = msg("Tuple")
    vt = msg("Valuet", tp)
    vt.oneof_decl.add().name = "oneof_values"
    _field(vt, "svalue", 1, "string", oneof=0)
    _field(vt, "ivalue", 2, "int32", oneof=0)
    _field(vt, "fvalue", 3, "float", oneof=0)
    _field(tp, "values", 1, "Tuple.Valuet", repeated=True)

    dc = msg("Dictionary")
    pair = msg("Pair", dc)
    _field(pair, "key", 1, "string")
    val = msg("Value", pair)
    val.oneof_decl.add().name = "oneof_values"
    _field(val, "svalue", 1, "string", oneof=0)
    _field(val, "ivalue", 2, "int32", oneof=0)
    _field(val, "fvalue", 3, "float", oneof=0)
    _field(val, "tvalue", 4, "Tuple", oneof=0)
    _field(val, "bvalue", 5, "bool", oneof=0)
    _field(pair, "value", 2, "Dictionary.Pair.Value")
    _field(dc, "pairs", 1, "Dictionary.Pair", repeated=True)

    dr = msg("DictRequest")
    _field(dr, "vocab", 1, "Dictionary")
    _field(dr, "client_id", 2, "int32")
    _field(dr, "nr_samples", 3, "int32")

    fu = msg("FeatureUnion")
    _field(fu, "dic", 1, "Dictionary", repeated=True)
    _field(fu, "initialNN", 2, "NNUpdate")
    _field(fu, "model_params", 3, "Dictionary")
    _field(fu, "model_type", 4, "string")

    mu = msg("ModelUpdate")
    for i, n in enumerate(MODEL_UPDATE_TENSOR_FIELDS, 1):
        _field(mu, n, i, "Tensor")
    _field(mu, "current_epoch", 25, "int32")
    _field(mu, "extra_tensors", 26, "Update", repeated=True)

    au = msg("AdamUpdate")
    st = msg("State", au)
    cs = msg("ContentState", st)
    _field(cs, "state_id", 1, "int64")
    _field(cs, "step", 2, "Tensor")
    _field(cs, "exp_avg", 3, "Tensor")
    _field(cs, "exp_avg_sq", 4, "Tensor")
    _field(st, "contentState", 2, "AdamUpdate.State.ContentState", repeated=True)
    _field(au, "state", 1, "AdamUpdate.State")
    pg = msg("ParamGroups", au)
    _field(pg, "lr", 1, "float")
    bt = msg("Betas", pg)
    _field(bt, "beta1", 1, "float")
    _field(bt, "beta2", 2, "float")
    _field(pg, "betas", 2, "AdamUpdate.ParamGroups.Betas")
    _field(pg, "eps", 3, "float")
    _field(pg, "weight_decay", 4, "float")
    _field(pg, "amsgrad", 5, "bool")
    _field(pg, "params", 6, "int32", repeated=True)
    _field(au, "paramGroups", 2, "AdamUpdate.ParamGroups")

    ou = msg("OptUpdate")
    ou.oneof_decl.add().name = "oneof_values"
    _field(ou, "adamUpdate", 1, "AdamUpdate", oneof=0)

    nn = msg("NNUpdate")
    _field(nn, "modelUpdate", 1, "ModelUpdate")
    _field(nn, "optUpdate", 2, "OptUpdate")

    _field(msg("ServerGetGradientRequest"), "iter", 1, "int64")

    def service(name, methods):
        s = fd.service.add()
        s.name = name
        for mname, inp, out in methods:
            m = s.method.add()
            m.name = mname
            m.input_type = "." + PACKAGE + "." + inp
            m.output_type = "." + PACKAGE + "." + out

    service("Federation", [("sendAggregatedTensor", "Empty", "ServerAggregatedTensorRequest"),
                           ("sendLocalDic", "DictRequest", "Reply"),
                           ("sendGlobalDicAndInitialNN", "Empty", "FeatureUnion"),
                           ("trainFederatedModel", "ClientTensorRequest", "Empty")])
    service("FederationServer", [("getGradient", "ServerGetGradientRequest", "ClientTensorRequest"),
                                 ("sendAggregatedTensor", "ServerAggregatedTensorRequest",
                                  "ClientReceivedResponse")])
    return fd


# the 24 tensor fields of ModelUpdate, in field-number order (1..24)
MODEL_UPDATE_TENSOR_FIELDS = (
    "prior_mean", "prior_variance", "beta", "topic_word_matrix",
    "inf_net_input_layer_weight", "inf_net_input_layer_bias",
    "inf_net_hiddens_l00_weight", "inf_net_hiddens_l_00_bias",
    "inf_net_f_mu_weight", "inf_net_f_mu_bias",
    "inf_net_f_mu_batchnorm_running_mean", "inf_net_f_mu_batchnorm_running_var",
    "inf_net_f_mu_batchnorm_num_batches_tracked",
    "inf_net_f_sigma_weight", "inf_net_f_sigma_bias",
    "inf_net_f_sigma_batchnorm_running_mean", "inf_net_f_sigma_batchnorm_running_var",
    "inf_net_f_sigma_batchnorm_num_batches_tracked",
    "beta_batchnorm_running_mean", "beta_batchnorm_running_var", "beta_batchnorm_num_batches_tracked",
    "best_components", "inf_net_adapt_bert_weight", "inf_net_adapt_bert_bias",
)

_POOL = _pool.DescriptorPool()
DESCRIPTOR = _POOL.Add(_build_file())


def _cls(name: str):
    return _mf.GetMessageClass(_POOL.FindMessageTypeByName(PACKAGE + "." + name))


Empty = _cls("Empty")
TensorShape = _cls("TensorShape")
Tensor = _cls("Tensor")
Update = _cls("Update")
MessageAdditionalData = _cls("MessageAdditionalData")
MessageHeader = _cls("MessageHeader")
ClientTensorRequest = _cls("ClientTensorRequest")
ServerAggregatedTensorRequest = _cls("ServerAggregatedTensorRequest")
ClientReceivedResponse = _cls("ClientReceivedResponse")
ServerReceivedResponse = _cls("ServerReceivedResponse")
Chunk = _cls("Chunk")
Request = _cls("Request")
Reply = _cls("Reply")
Tuple = _cls("Tuple")
Dictionary = _cls("Dictionary")
DictRequest = _cls("DictRequest")
FeatureUnion = _cls("FeatureUnion")
ModelUpdate = _cls("ModelUpdate")
AdamUpdate = _cls("AdamUpdate")
OptUpdate = _cls("OptUpdate")
NNUpdate = _cls("NNUpdate")
ServerGetGradientRequest = _cls("ServerGetGradientRequest")


class _MessageTypeEnum:
    """``federated_pb2.MessageType.CLIENT_TENSOR_SEND`` style access plus Name()/Value()."""

    _desc = _POOL.FindEnumTypeByName(PACKAGE + ".MessageType")

    def __init__(self) -> None:
        for v in self._desc.values:
            setattr(self, v.name, v.number)

    def Name(self, number: int) -> str:
        return self._desc.values_by_number[number].name

    def Value(self, name: str) -> int:
        return self._desc.values_by_name[name].number


MessageType = _MessageTypeEnum()
for _v in MessageType._desc.values:  # module-level constants like generated code
    globals()[_v.name] = _v.number
