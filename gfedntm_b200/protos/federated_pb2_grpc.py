"""gRPC client stubs / servicer bases for the two services of ``federated.proto``.

Written by hand on top of ``grpc``'s generic handlers (no ``grpc_tools`` code generation); the
method paths (``/federated.Federation/...``) and (de)serialisers match generated stubs.
"""
from __future__ import annotations

import grpc

from . import federated_pb2 as pb

_FEDERATION = {
    "sendAggregatedTensor": (pb.Empty, pb.ServerAggregatedTensorRequest),
    "sendLocalDic": (pb.DictRequest, pb.Reply),
    "sendGlobalDicAndInitialNN": (pb.Empty, pb.FeatureUnion),
    "trainFederatedModel": (pb.ClientTensorRequest, pb.Empty),
}
_FEDERATION_SERVER = {
    "getGradient": (pb.ServerGetGradientRequest, pb.ClientTensorRequest),
    "sendAggregatedTensor": (pb.ServerAggregatedTensorRequest, pb.ClientReceivedResponse),
}


def _make_stub(service: str, table: dict):
    class _Stub:
        def __init__(self, channel: grpc.Channel) -> None:
            for name, (req, resp) in table.items():
                setattr(self, name, channel.unary_unary(f"/federated.{service}/{name}",
                                                        request_serializer=req.SerializeToString,
                                                        response_deserializer=resp.FromString))
    _Stub.__name__ = service + "Stub"
    _Stub.__qualname__ = _Stub.__name__
    return _Stub


def _make_servicer(service: str, table: dict):
    def _unimplemented(self, request, context):
        context.set_code(grpc.StatusCode.UNIMPLEMENTED)
        context.set_details("Method not implemented!")
        raise NotImplementedError("Method not implemented!")

    cls = type(service + "Servicer", (object,), {name: _unimplemented for name in table})
    return cls


def _make_adder(service: str, table: dict):
    def add(servicer, server: grpc.Server) -> None:
        handlers = {
            name: grpc.unary_unary_rpc_method_handler(getattr(servicer, name),
                                                      request_deserializer=req.FromString,
                                                      response_serializer=resp.SerializeToString)
            for name, (req, resp) in table.items()}
        server.add_generic_rpc_handlers((grpc.method_handlers_generic_handler(f"federated.{service}", handlers),))
    return add


FederationStub = _make_stub("Federation", _FEDERATION)
FederationServicer = _make_servicer("Federation", _FEDERATION)
add_FederationServicer_to_server = _make_adder("Federation", _FEDERATION)

FederationServerStub = _make_stub("FederationServer", _FEDERATION_SERVER)
FederationServerServicer = _make_servicer("FederationServer", _FEDERATION_SERVER)
add_FederationServerServicer_to_server = _make_adder("FederationServer", _FEDERATION_SERVER)
