"""Wire helpers of the actor runtime: value descriptors, (de)serialisation, RPC client.

A *descriptor* is what travels through the control plane instead of the value:
``("inline", pickle_bytes)``, ``("shm", segment_name, kind)`` or ``("error", pickle_bytes)``.
Bulk values (Arrow tables, large bytes/pickles) are written once into the shared-memory
store by whoever produced them and never pass through a socket -- the replacement for
Ray's plasma hand-off used by reference ``python/raydp/spark/dataset.py:78-84``.
"""
from __future__ import annotations

import itertools
import os
import pickle
import threading
import uuid
from multiprocessing.connection import Client as _MPClient
from typing import Any, Dict, Optional, Tuple

import cloudpickle
import pyarrow as pa

from raydp_b200 import store

INLINE_LIMIT = 256 * 1024

ENV_HEAD = "RAYDP_B200_HEAD"
ENV_AUTH = "RAYDP_B200_AUTH"
ENV_NODE = "RAYDP_B200_NODE"
ENV_NODE_IP = "RAYDP_B200_NODE_IP"
ENV_SESSION = "RAYDP_B200_SESSION"
ENV_WORKER = "RAYDP_B200_WORKER_ID"
ENV_NAMESPACE = "RAYDP_B200_NAMESPACE"


class RuntimeErrorBase(Exception):
    pass


class TaskError(RuntimeErrorBase):
    """A remote function or actor method raised; carries the remote traceback text."""

    def __init__(self, message: str, cause: Optional[BaseException] = None):
        super().__init__(message)
        self.cause = cause


class ActorDiedError(RuntimeErrorBase):
    pass


class GetTimeoutError(RuntimeErrorBase, TimeoutError):
    pass


class ObjectLostError(RuntimeErrorBase):
    pass


def new_id() -> str:
    return uuid.uuid4().hex


def segment_name(session: str, node_id: str, oid: str) -> str:
    return f"rdp_{session}_{node_id}_{oid}"


def dumps(value: Any) -> bytes:
    return cloudpickle.dumps(value, protocol=5)


def loads(blob: bytes) -> Any:
    return pickle.loads(blob)


def make_descriptor(value: Any, session: str, node_id: str, oid: str) -> Tuple:
    """Serialise ``value``; Arrow and large payloads go to shared memory."""
    if isinstance(value, (pa.Table, pa.RecordBatch)):
        name = segment_name(session, node_id, oid)
        store.put_arrow(name, value)
        return ("shm", name, store.KIND_ARROW)
    if isinstance(value, (bytes, bytearray, memoryview)) and len(value) > INLINE_LIMIT:
        name = segment_name(session, node_id, oid)
        store.put_bytes(name, value)
        return ("shm", name, store.KIND_BYTES)
    blob = dumps(value)
    if len(blob) > INLINE_LIMIT:
        name = segment_name(session, node_id, oid)
        store.put_bytes(name, blob, kind=store.KIND_PICKLE)
        return ("shm", name, store.KIND_PICKLE)
    return ("inline", blob)


def error_descriptor(exc: BaseException, tb_text: str = "") -> Tuple:
    try:
        blob = dumps((exc, tb_text))
    except Exception:  # noqa: BLE001 - unpicklable exception
        blob = dumps((RuntimeError(repr(exc)), tb_text))
    return ("error", blob)


def read_descriptor(desc: Tuple, pin: bool = False) -> Any:
    kind = desc[0]
    if kind == "inline":
        return loads(desc[1])
    if kind == "shm":
        try:
            return store.get(desc[1], pin=pin)
        except store.ObjectLostError as exc:
            raise ObjectLostError(str(exc)) from None
    if kind == "error":
        exc, tb_text = loads(desc[1])
        if isinstance(exc, RuntimeErrorBase):
            raise exc
        raise TaskError(f"{type(exc).__name__}: {exc}\n{tb_text}", cause=exc)
    raise ValueError(f"bad descriptor {desc!r}")


class RpcClient:
    """Request/response multiplexing over one ``multiprocessing.connection``."""

    def __init__(self, address: str, authkey: bytes):
        self._conn = _MPClient(address, family="AF_UNIX", authkey=authkey)
        self._send_lock = threading.Lock()
        self._pending: Dict[int, list] = {}
        self._pending_lock = threading.Lock()
        self._ids = itertools.count(1)
        self._closed = False
        self._reader = threading.Thread(target=self._read_loop, name="rdp-rpc-reader", daemon=True)
        self._reader.start()

    def _read_loop(self) -> None:
        try:
            while True:
                req_id, ok, value = self._conn.recv()
                with self._pending_lock:
                    slot = self._pending.pop(req_id, None)
                if slot is not None:
                    slot[1] = (ok, value)
                    slot[0].set()
        except (EOFError, OSError, TypeError):
            pass
        finally:
            self._closed = True
            with self._pending_lock:
                slots = list(self._pending.values())
                self._pending.clear()
            for slot in slots:
                slot[1] = (False, dumps(RuntimeErrorBase("runtime head connection closed")))
                slot[0].set()

    def request(self, op: str, **kwargs) -> Any:
        if self._closed:
            raise RuntimeErrorBase("runtime head connection closed")
        req_id = next(self._ids)
        slot = [threading.Event(), None]
        with self._pending_lock:
            self._pending[req_id] = slot
        with self._send_lock:
            self._conn.send((req_id, op, kwargs))
        slot[0].wait()
        ok, value = slot[1]
        if ok:
            return value
        exc = loads(value) if isinstance(value, (bytes, bytearray)) else value
        raise exc

    def close(self) -> None:
        self._closed = True
        try:
            self._conn.close()
        except OSError:
            pass


def env_for_worker(head_addr: str, authkey: bytes, node_id: str, node_ip: str, session: str,
                   worker_id: str, namespace: str) -> Dict[str, str]:
    env = dict(os.environ)
    env[ENV_HEAD] = head_addr
    env[ENV_AUTH] = authkey.hex()
    env[ENV_NODE] = node_id
    env[ENV_NODE_IP] = node_ip
    env[ENV_SESSION] = session
    env[ENV_WORKER] = worker_id
    env[ENV_NAMESPACE] = namespace
    return env
