"""Python face of the C++ shared-memory object store (``csrc/store/store.cc``).

Parity: the role Ray's plasma store plays for the reference -- ``ray.put``/``ray.get`` of
``pa.Table`` blocks and Arrow IPC ``bytes`` (reference ``python/raydp/spark/dataset.py:78-84``,
``:222``, ``:307-311``).  Values are written once (Arrow IPC stream, raw bytes or a pickle)
into a sealed segment and read back zero-copy: the returned ``pa.Table``'s buffers point
into the mapping, which stays alive as long as any of them does.
"""
from __future__ import annotations

import ctypes
import errno
import json
import os
import pickle
import time
from typing import Any, Optional, Tuple

import pyarrow as pa

from raydp_b200._native import store_lib

KIND_ARROW = "arrow"
KIND_BYTES = "bytes"
KIND_PICKLE = "pickle"
_META_RESERVE = 256


class StoreError(RuntimeError):
    pass


class ObjectLostError(StoreError):
    """The segment backing an object is gone (its node was removed / data was cleaned up)."""


pin_failures = 0       # objects whose mapping could not be page-locked (CUDA absent, or the registration failed)
pinned_objects = 0


class Mapping:
    """Owns one mapped segment; closing unmaps (and unpins) it."""

    __slots__ = ("_h", "name", "__weakref__")

    def __init__(self, handle: int, name: str):
        self._h = ctypes.c_void_p(handle)
        self.name = name

    @property
    def data_address(self) -> int:
        return store_lib().rdp_store_data(self._h)

    @property
    def data_size(self) -> int:
        return store_lib().rdp_store_data_size(self._h)

    def meta(self) -> dict:
        lib = store_lib()
        size = lib.rdp_store_meta_size(self._h)
        raw = ctypes.string_at(lib.rdp_store_meta(self._h), size)
        return json.loads(raw.rstrip(b"\0").decode() or "{}")

    def buffer(self) -> pa.Buffer:
        return pa.foreign_buffer(self.data_address, self.data_size, base=self)

    def mutable_buffer(self) -> pa.Buffer:
        raw = (ctypes.c_char * max(1, self.data_size)).from_address(self.data_address)
        return pa.py_buffer(memoryview(raw))

    def pin(self) -> bool:
        """cudaHostRegister this mapping; False when CUDA is absent or registration fails."""
        return store_lib().rdp_store_pin(self._h) == 0

    def close(self) -> None:
        if self._h:
            store_lib().rdp_store_close(self._h)
            self._h = ctypes.c_void_p(0)

    def __del__(self):  # noqa: D105
        try:
            self.close()
        except Exception:  # noqa: BLE001 - interpreter shutdown
            pass


def _check(rc: int, what: str, name: str) -> None:
    if rc < 0:
        code = -rc
        if code == errno.ENOENT:
            raise ObjectLostError(f"{what}: object {name!r} is not in the store (lost or freed)")
        raise StoreError(f"{what} {name!r} failed: {os.strerror(code) if code < 1000 else code}")


def _create(name: str, data_size: int, meta: dict) -> Mapping:
    lib = store_lib()
    blob = json.dumps(meta).encode()
    handle = ctypes.c_void_p()
    rc = lib.rdp_store_create(name.encode(), int(data_size), max(len(blob), _META_RESERVE),
                              ctypes.byref(handle))
    if rc == -errno.EEXIST:
        # leftover of a producer that died mid-write (its retry reuses the object id)
        lib.rdp_store_unlink(name.encode())
        rc = lib.rdp_store_create(name.encode(), int(data_size), max(len(blob), _META_RESERVE),
                                  ctypes.byref(handle))
    _check(rc, "create", name)
    ctypes.memmove(lib.rdp_store_meta(handle), blob, len(blob))
    return Mapping(handle.value, name)


def put_arrow(name: str, table: pa.Table, extra_meta: Optional[dict] = None) -> int:
    """Write ``table`` as one Arrow IPC stream (8-byte aligned buffers); returns bytes used."""
    if isinstance(table, pa.RecordBatch):
        table = pa.Table.from_batches([table])
    opts = pa.ipc.IpcWriteOptions()
    sink = pa.MockOutputStream()
    with pa.ipc.new_stream(sink, table.schema, options=opts) as writer:
        writer.write_table(table)
    size = sink.size()
    meta = {"kind": KIND_ARROW, "num_rows": table.num_rows, "nbytes": int(table.nbytes)}
    if extra_meta:
        meta.update(extra_meta)
    mapping = _create(name, size, meta)
    out = pa.FixedSizeBufferWriter(mapping.mutable_buffer())
    out.set_memcopy_threads(4)
    with pa.ipc.new_stream(out, table.schema, options=opts) as writer:
        writer.write_table(table)
    store_lib().rdp_store_seal(mapping._h)
    mapping.close()
    return size


def put_bytes(name: str, payload, kind: str = KIND_BYTES, extra_meta: Optional[dict] = None) -> int:
    view = memoryview(payload).cast("B")
    meta = {"kind": kind, "nbytes": view.nbytes}
    if extra_meta:
        meta.update(extra_meta)
    mapping = _create(name, view.nbytes, meta)
    if view.nbytes:
        src = (ctypes.c_char * view.nbytes).from_buffer_copy(view) if view.readonly else \
            (ctypes.c_char * view.nbytes).from_buffer(view)
        store_lib().rdp_store_parallel_copy(mapping.data_address, ctypes.addressof(src),
                                            view.nbytes, 4)
    store_lib().rdp_store_seal(mapping._h)
    mapping.close()
    return view.nbytes


def put_pickle(name: str, value: Any) -> int:
    import cloudpickle
    return put_bytes(name, cloudpickle.dumps(value, protocol=5), kind=KIND_PICKLE)


def open_mapping(name: str, writable: bool = False, wait_s: float = 5.0) -> Mapping:
    lib = store_lib()
    handle = ctypes.c_void_p()
    deadline = time.time() + wait_s
    while True:
        rc = lib.rdp_store_open(name.encode(), 1 if writable else 0, ctypes.byref(handle))
        if rc == -errno.EAGAIN and time.time() < deadline:
            time.sleep(0.001)
            continue
        _check(rc, "open", name)
        return Mapping(handle.value, name)


def get(name: str, pin: bool = False) -> Any:
    """Read an object back.  Arrow tables and bytes are zero-copy views of shared memory."""
    value, _ = get_with_mapping(name, pin=pin)
    return value


def get_with_mapping(name: str, pin: bool = False) -> Tuple[Any, Mapping]:
    mapping = open_mapping(name, writable=pin)
    if pin and not mapping.pin():
        # pageable memory still works with cudaMemcpyAsync, but the copy is staged by the driver and
        # blocks the calling thread: say so once instead of silently losing the DMA overlap
        global pin_failures
        pin_failures += 1
        if pin_failures == 1:
            import logging
            logging.getLogger(__name__).warning(
                "cudaHostRegister failed for object %s: host->device copies from it will not be asynchronous", name)
    elif pin:
        global pinned_objects
        pinned_objects += 1
    meta = mapping.meta()
    kind = meta.get("kind", KIND_BYTES)
    buf = mapping.buffer()
    if kind == KIND_ARROW:
        return pa.ipc.open_stream(buf).read_all(), mapping
    if kind == KIND_PICKLE:
        return pickle.loads(memoryview(buf)), mapping
    return buf.to_pybytes() if meta.get("copy", True) else buf, mapping


def exists(name: str) -> bool:
    return bool(store_lib().rdp_store_exists(name.encode()))


def delete(name: str) -> bool:
    return store_lib().rdp_store_unlink(name.encode()) == 0


def delete_prefix(prefix: str) -> int:
    return int(store_lib().rdp_store_unlink_prefix(prefix.encode()))


def usage(prefix: str) -> Tuple[int, int]:
    """``(object_count, bytes)`` of everything under ``prefix``."""
    nbytes = ctypes.c_uint64()
    count = store_lib().rdp_store_usage(prefix.encode(), ctypes.byref(nbytes))
    return int(count), int(nbytes.value)
