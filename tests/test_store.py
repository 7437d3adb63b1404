"""C++ shared-memory object store: round trips, zero copy, loss."""
import uuid

import numpy as np
import pyarrow as pa
import pytest

from raydp_b200 import store


def _name():
    return "rdp_test_" + uuid.uuid4().hex


def test_arrow_roundtrip_is_zero_copy():
    name = _name()
    table = pa.table({"a": np.arange(100000, dtype=np.int64), "b": np.random.rand(100000)})
    used = store.put_arrow(name, table)
    assert used >= table.nbytes
    try:
        got, mapping = store.get_with_mapping(name)
        assert got.equals(table)
        addr = got.column("a").chunk(0).buffers()[1].address
        assert mapping.data_address <= addr < mapping.data_address + mapping.data_size
        assert mapping.meta()["num_rows"] == 100000
    finally:
        assert store.delete(name)
    # the mapping keeps the pages alive after unlink
    assert got.column("a")[99999].as_py() == 99999


def test_bytes_and_pickle_roundtrip():
    n1, n2 = _name(), _name()
    payload = bytes(range(256)) * 4096
    store.put_bytes(n1, payload)
    store.put_pickle(n2, {"k": [1, 2, 3]})
    try:
        assert store.get(n1) == payload
        assert store.get(n2) == {"k": [1, 2, 3]}
    finally:
        store.delete(n1)
        store.delete(n2)


def test_missing_object_raises_lost():
    with pytest.raises(store.ObjectLostError):
        store.get(_name())


def test_prefix_delete_and_usage():
    prefix = "rdp_test_" + uuid.uuid4().hex[:8] + "_"
    for i in range(3):
        store.put_bytes(prefix + str(i), b"x" * 1000)
    count, nbytes = store.usage(prefix)
    assert count == 3 and nbytes >= 3000
    assert store.delete_prefix(prefix) == 3
    assert store.usage(prefix)[0] == 0


def test_recreate_after_crashed_producer():
    name = _name()
    store.put_bytes(name, b"first")
    store.put_bytes(name, b"second")      # EEXIST path: stale segment is replaced
    assert store.get(name) == b"second"
    store.delete(name)
