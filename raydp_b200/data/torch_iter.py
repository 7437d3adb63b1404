"""CPU batch iterator behind ``Dataset.to_torch`` plus a background prefetcher.

Parity: ``dataset.to_torch(...)`` as consumed by reference ``torch/estimator.py:226-241`` and the
``PrefetchedDataLoader`` capability of reference ``torch/torch_ml_dataset.py:69-111``
(background thread + bounded queue).  Columns are converted block-wise Arrow -> numpy without
pandas; the GPU ingestion path (``raydp_b200.ops.ingest``) replaces this on CUDA workers.
"""
from __future__ import annotations

import queue
import threading
from typing import Any, Iterator, List, Optional, Sequence

import numpy as np
import pyarrow as pa


def _column_numpy(table: pa.Table, name: str, null_fill: float = 0.0, floating: bool = False) -> np.ndarray:
    """One column as a numpy array.  NULLs: ``null_fill`` when the consumer wants a floating tensor
    (``floating=True``; NaN reproduces the reference's pandas path), else 0."""
    col = table.column(name)
    arr = col.combine_chunks() if isinstance(col, pa.ChunkedArray) else col
    if pa.types.is_timestamp(arr.type) or pa.types.is_date(arr.type):
        arr = arr.cast(pa.int64())
    if pa.types.is_boolean(arr.type):
        arr = arr.cast(pa.int8())
    if arr.null_count:
        if floating:
            mask = arr.is_null().to_numpy(zero_copy_only=False)
            out = arr.fill_null(0).to_numpy(zero_copy_only=False).astype(np.float64, copy=True)
            out[mask] = null_fill
            return out
        arr = arr.fill_null(0)
    return arr.to_numpy(zero_copy_only=False)


def _is_vector(table: pa.Table, name: str) -> bool:
    t = table.schema.field(name).type
    return pa.types.is_list(t) or pa.types.is_large_list(t) or pa.types.is_fixed_size_list(t)


def _column_matrix(table: pa.Table, name: str) -> np.ndarray:
    """A fixed-width ``array<number>`` column (what ``VectorAssembler`` produces) as an ``[n, width]`` array."""
    import pyarrow.compute as pc
    col = table.column(name)
    arr = col.combine_chunks() if isinstance(col, pa.ChunkedArray) else col
    n = len(arr)
    if n == 0:
        return np.zeros((0, 0))
    lengths = pc.list_value_length(arr)
    width = pc.max(lengths).as_py() or 0
    if arr.null_count or pc.min(lengths).as_py() != width:
        raise ValueError(f"feature column {name!r} must hold vectors of one length without nulls")
    flat = pc.list_flatten(arr)
    if flat.null_count:
        flat = flat.fill_null(0)
    return flat.to_numpy(zero_copy_only=False).reshape(n, width)


def _np_dtype(torch_dtype):
    import torch
    if torch_dtype is None:
        return None
    return {torch.float32: np.float32, torch.float64: np.float64, torch.float16: np.float16,
            torch.int64: np.int64, torch.int32: np.int32, torch.int16: np.int16,
            torch.int8: np.int8, torch.uint8: np.uint8, torch.bool: np.bool_,
            torch.bfloat16: np.float32}.get(torch_dtype, None)


def _is_float_np(dt) -> bool:
    return dt is not None and np.issubdtype(dt, np.floating)


def block_to_arrays(table: pa.Table, feature_columns: Sequence[str], feature_dtype,
                    label_column: Optional[str], label_dtype, null_fill: float = float("nan")):
    """One Arrow block -> ``([n, F] features, [n, 1] label)`` numpy arrays."""
    f_np = _np_dtype(feature_dtype) or np.float32
    ffl = _is_float_np(f_np)
    if any(_is_vector(table, name) for name in feature_columns):
        # vector columns contribute their width: [scalar | vector(k) | ...] -> [n, 1 + k + ...]
        parts = [_column_matrix(table, name) if _is_vector(table, name)
                 else _column_numpy(table, name, null_fill, ffl).reshape(-1, 1) for name in feature_columns]
        feats = np.concatenate([p.astype(f_np, copy=False) for p in parts], axis=1) if table.num_rows else \
            np.empty((0, len(feature_columns)), dtype=f_np)
    else:
        feats = np.empty((table.num_rows, len(feature_columns)), dtype=f_np)
        for j, name in enumerate(feature_columns):
            feats[:, j] = _column_numpy(table, name, null_fill, ffl)
    label = None
    if label_column is not None:
        l_np = _np_dtype(label_dtype) or np.float32
        label = _column_numpy(table, label_column, null_fill, _is_float_np(l_np)).astype(l_np, copy=False).reshape(-1, 1)
    return feats, label


class TorchBatchIterable:
    def __init__(self, dataset, label_column, feature_columns, label_column_dtype,
                 feature_column_dtypes, batch_size: int, drop_last: bool,
                 unsqueeze_label: bool = True, null_fill: float = float("nan")):
        import torch
        self._ds = dataset
        self._label = label_column
        names = dataset.columns()
        if feature_columns is None:
            feature_columns = [n for n in names if n != label_column]
        self._features: List[str] = list(feature_columns)
        self._label_dtype = label_column_dtype
        self._multi = isinstance(feature_column_dtypes, (list, tuple))
        if self._multi and len(feature_column_dtypes) != len(self._features):
            raise ValueError("feature_column_dtypes must have one dtype per feature column")
        self._feature_dtypes = feature_column_dtypes
        self._batch = int(batch_size)
        self._drop_last = drop_last
        self._unsqueeze = unsqueeze_label
        self._null_fill = float(null_fill)
        self._torch = torch

    def __len__(self) -> int:
        n = self._ds.count()
        return n // self._batch if self._drop_last else -(-n // self._batch)

    def _convert(self, table: pa.Table):
        torch = self._torch
        if self._multi:
            feats = []
            for name, dt in zip(self._features, self._feature_dtypes):
                if _is_vector(table, name):
                    t = torch.as_tensor(np.ascontiguousarray(_column_matrix(table, name)))
                else:
                    fl = dt is None or dt.is_floating_point
                    t = torch.as_tensor(np.ascontiguousarray(
                        _column_numpy(table, name, self._null_fill, fl))).reshape(-1, 1)
                feats.append(t.to(dt) if dt is not None else t)
            label = None
            if self._label is not None:
                lfl = self._label_dtype is None or self._label_dtype.is_floating_point
                label = torch.as_tensor(np.ascontiguousarray(
                    _column_numpy(table, self._label, self._null_fill, lfl)))
                label = label.to(self._label_dtype) if self._label_dtype is not None else label
                label = label.reshape(-1, 1) if self._unsqueeze else label
            return feats, label
        f, l = block_to_arrays(table, self._features, self._feature_dtypes, self._label,
                               self._label_dtype, self._null_fill)
        ft = torch.from_numpy(f)
        if self._feature_dtypes is not None and ft.dtype != self._feature_dtypes:
            ft = ft.to(self._feature_dtypes)
        lt = None
        if l is not None:
            lt = torch.from_numpy(np.ascontiguousarray(l))
            if self._label_dtype is not None and lt.dtype != self._label_dtype:
                lt = lt.to(self._label_dtype)
            if not self._unsqueeze:
                lt = lt.reshape(-1)
        return ft, lt

    def __iter__(self) -> Iterator[Any]:
        """Blocks are converted to tensors ONCE (one numpy conversion per column and block) and the batches are
        slices of those tensors; a batch that straddles two blocks is stitched from the tail of one and the head of
        the next.  (Converting per batch costs ~20 small Arrow -> numpy conversions for every 64 rows: the CPU
        estimator spent more time there than in the model.)"""
        torch = self._torch
        cols = self._features + ([self._label] if self._label is not None else [])
        B = self._batch

        def rows_of(item):
            feats = item[0]
            return (feats[0] if self._multi else feats).shape[0]

        def cut(item, lo, hi):
            feats, label = item
            f = [t[lo:hi] for t in feats] if self._multi else feats[lo:hi]
            return f, (label[lo:hi] if label is not None else None)

        def join(a, b):
            fa, la = a
            fb, lb = b
            f = [torch.cat([x, y]) for x, y in zip(fa, fb)] if self._multi else torch.cat([fa, fb])
            return f, (torch.cat([la, lb]) if la is not None else None)

        carry = None
        for table in self._ds.iter_tables():
            if table.num_rows == 0:
                continue
            item = self._convert(table.select(cols))
            n, off = rows_of(item), 0
            if carry is not None:
                need = B - rows_of(carry)
                take = min(need, n)
                carry = join(carry, cut(item, 0, take))
                off = take
                if rows_of(carry) < B:
                    continue
                yield carry
                carry = None
            while n - off >= B:
                yield cut(item, off, off + B)
                off += B
            if off < n:
                carry = cut(item, off, n)
        if carry is not None and not self._drop_last:
            yield carry


class PrefetchedLoader:
    """Wrap any iterable; a daemon thread keeps up to ``max_size`` items ready."""

    _END = object()

    def __init__(self, base, max_size: int = 5):
        self._base = base
        self._max = max_size

    def __len__(self):
        return len(self._base)

    def __iter__(self):
        q: "queue.Queue" = queue.Queue(maxsize=self._max)
        stop = threading.Event()

        def work():
            try:
                for item in self._base:
                    if stop.is_set():
                        return
                    q.put(item)
                q.put(self._END)
            except BaseException as exc:  # noqa: BLE001
                q.put(exc)
        thread = threading.Thread(target=work, name="rdp-prefetch", daemon=True)
        thread.start()
        try:
            while True:
                item = q.get()
                if item is self._END:
                    return
                if isinstance(item, BaseException):
                    raise item
                yield item
        finally:
            stop.set()
