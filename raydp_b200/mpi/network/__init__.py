"""gRPC plumbing of the MPI launcher without generated stubs.

Parity: reference ``python/raydp/mpi/network/network.proto:19-83`` defines
``DriverService{RegisterWorker, RegisterWorkerService, RegisterFuncResult}`` and
``WorkerService{RunFunction, Stop}``.  The same five RPCs are served here through gRPC *generic
handlers*; messages are small dicts serialised with pickle (the payload that matters -- the
cloudpickled function / result -- is opaque bytes either way), so no protoc / grpc_tools step is
needed at build time.
"""
import pickle
from typing import Callable, Dict

import grpc

from raydp_b200.mpi.constants import MAX_MESSAGE_BYTES

DRIVER_SERVICE = "raydp.mpi.DriverService"
WORKER_SERVICE = "raydp.mpi.WorkerService"
OPTIONS = [("grpc.max_send_message_length", MAX_MESSAGE_BYTES),
           ("grpc.max_receive_message_length", MAX_MESSAGE_BYTES),
           ("grpc.enable_http_proxy", 0)]


def _ser(obj) -> bytes:
    return pickle.dumps(obj, protocol=5)


def _de(blob: bytes):
    return pickle.loads(blob)


def generic_handler(service: str, methods: Dict[str, Callable]):
    handlers = {name: grpc.unary_unary_rpc_method_handler(
        (lambda fn: (lambda request, context: fn(request)))(fn),
        request_deserializer=_de, response_serializer=_ser) for name, fn in methods.items()}
    return grpc.method_handlers_generic_handler(service, handlers)


class Stub:
    """Callable attributes for each RPC of ``service`` on ``channel``."""

    def __init__(self, channel: grpc.Channel, service: str, methods):
        for name in methods:
            setattr(self, name, channel.unary_unary(f"/{service}/{name}", request_serializer=_ser,
                                                    response_deserializer=_de))


def driver_stub(channel) -> Stub:
    return Stub(channel, DRIVER_SERVICE, ["RegisterWorker", "RegisterWorkerService", "RegisterFuncResult"])


def worker_stub(channel) -> Stub:
    return Stub(channel, WORKER_SERVICE, ["RunFunction", "Stop"])


def create_insecure_channel(address: str) -> grpc.Channel:
    """Channel with the http proxy disabled (reference ``mpi/utils.py:87-96``)."""
    return grpc.insecure_channel(address, options=OPTIONS)
