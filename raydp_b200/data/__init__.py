"""Distributed dataset of Arrow blocks in the object store (the ``ray.data`` surface RayDP uses).

Parity: ``from_arrow_refs`` (reference ``spark/dataset.py:222``),
``from_items / from_arrow / from_pandas / read_parquet``, ``count / take / show / schema``
(schema exposes ``.base_schema``, reference ``spark/dataset.py:298-299``),
``iter_internal_ref_bundles`` (reference ``test_spark_cluster.py:262-299``), ``split``,
``split_at_indices``, ``random_shuffle`` (reference ``torch/estimator.py:335-338``),
``to_torch(feature_columns, feature_column_dtypes, label_column, label_column_dtype,
batch_size, drop_last)`` (reference ``torch/estimator.py:226-241``) and ``to_spark``
(reference ``README.md:117``).

A block is an ObjectRef to a ``pa.Table`` plus an optional row window, so equal-row splits for
data-parallel ranks are zero-copy slices rather than rewritten blocks.
"""
from __future__ import annotations

import builtins
import math
import os
from dataclasses import dataclass
from typing import Any, Callable, Dict, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np
import pyarrow as pa

from raydp_b200 import runtime

builtins_range = builtins.range

__all__ = ["Dataset", "BlockMetadata", "from_arrow_refs", "from_arrow", "from_items",
           "from_pandas", "from_numpy", "from_spark", "read_parquet", "read_csv", "range"]


@dataclass
class BlockMetadata:
    num_rows: int
    size_bytes: int
    schema: Optional[pa.Schema] = None
    input_files: Optional[List[str]] = None
    exec_stats: Optional[Any] = None


@dataclass
class Block:
    ref: Any                 # ObjectRef -> pa.Table
    offset: int              # first row of this view inside the stored table
    num_rows: int            # rows in this view
    stored_rows: int         # rows of the stored table
    size_bytes: int = 0

    def load(self, pin: bool = False) -> pa.Table:
        table = runtime.get(self.ref, _pin=pin)
        if isinstance(table, (bytes, memoryview, pa.Buffer)):
            table = pa.ipc.open_stream(table).read_all()
        if self.offset or self.num_rows != table.num_rows:
            table = table.slice(self.offset, self.num_rows)
        return table


class RefBundle:
    def __init__(self, blocks: List[Tuple[Any, BlockMetadata]]):
        self.blocks = blocks
        self.block_refs = [b[0] for b in blocks]

    def num_rows(self) -> int:
        return sum(m.num_rows for _, m in self.blocks)


class Schema:
    """``ds.schema()`` result: wraps the Arrow schema like ``ray.data.Schema``."""

    def __init__(self, base_schema: pa.Schema):
        self.base_schema = base_schema

    @property
    def names(self) -> List[str]:
        return list(self.base_schema.names)

    @property
    def types(self) -> List[pa.DataType]:
        return list(self.base_schema.types)

    def __repr__(self):
        return "Column  Type\n------  ----\n" + "\n".join(
            f"{n}  {t}" for n, t in zip(self.names, self.types))

    def __eq__(self, other):
        return isinstance(other, Schema) and other.base_schema == self.base_schema


def _block_stats(table: pa.Table) -> Tuple[int, int]:
    return table.num_rows, int(table.nbytes)


@runtime.remote(num_cpus=0)
def _stat_block(table) -> Optional[Tuple[int, int, bytes]]:
    if table is None:
        return None
    if not isinstance(table, pa.Table):
        table = pa.ipc.open_stream(table).read_all()
    return table.num_rows, int(table.nbytes), table.schema.serialize().to_pybytes()


@runtime.remote(num_cpus=0)
def _shuffle_map(table: pa.Table, offset: int, num_rows: int, n_out: int, seed: int, idx: int):
    table = table.slice(offset, num_rows)
    rng = np.random.default_rng([seed, idx])
    dest = rng.integers(0, n_out, size=table.num_rows)
    out = []
    for d in builtins_range(n_out):
        piece = table.filter(pa.array(dest == d))
        out.append(runtime.put(piece.combine_chunks()) if piece.num_rows else None)
    return out


@runtime.remote(num_cpus=0)
def _shuffle_reduce(seed: int, idx: int, *pieces):
    tables = [p for p in pieces if p is not None and p.num_rows]
    if not tables:
        return None
    table = pa.concat_tables(tables).combine_chunks()
    perm = np.random.default_rng([seed, idx, 1]).permutation(table.num_rows)
    return table.take(pa.array(perm))


@runtime.remote(num_cpus=0)
def _map_block(fn, table: pa.Table, offset: int, num_rows: int, batch_format: str):
    table = table.slice(offset, num_rows)
    if batch_format == "pandas":
        out = fn(table.to_pandas())
        return pa.Table.from_pandas(out, preserve_index=False)
    if batch_format == "numpy":
        out = fn({n: table.column(n).to_numpy(zero_copy_only=False) for n in table.column_names})
        return pa.table({k: pa.array(v) for k, v in out.items()})
    out = fn(table)
    return out if isinstance(out, pa.Table) else pa.Table.from_pandas(out, preserve_index=False)


@runtime.remote(num_cpus=0)
def _read_parquet_file(path: str, columns: Optional[List[str]]):
    import pyarrow.parquet as pq
    return pq.read_table(path, columns=columns)


@runtime.remote(num_cpus=0)
def _read_csv_file(path: str):
    import pyarrow.csv as pacsv
    return pacsv.read_csv(path)


def _in_waves(submit: Callable[[int], Any], n: int, width: Optional[int] = None) -> List[Any]:
    """Run ``n`` zero-CPU tasks at most ``width`` at a time (bounded worker processes)."""
    width = width or max(2, min(8, os.cpu_count() or 2))
    refs: List[Any] = [None] * n
    pending: List[Tuple[int, Any]] = []
    nxt = 0
    while nxt < n or pending:
        while nxt < n and len(pending) < width:
            refs[nxt] = submit(nxt)
            pending.append((nxt, refs[nxt]))
            nxt += 1
        ready, _ = runtime.wait([r for _, r in pending], num_returns=1, timeout=None)
        done = set(r.hex() for r in ready)
        pending = [(i, r) for i, r in pending if r.hex() not in done]
    return refs


class Dataset:
    def __init__(self, blocks: List[Block], schema: Optional[pa.Schema]):
        self._blocks = blocks
        self._schema = schema

    # ------------------------------------------------------------------ metadata
    def num_blocks(self) -> int:
        return len(self._blocks)

    def count(self) -> int:
        return sum(b.num_rows for b in self._blocks)

    def size_bytes(self) -> int:
        return sum(b.size_bytes for b in self._blocks)

    def schema(self) -> Optional[Schema]:
        if self._schema is None and self._blocks:
            self._schema = self._blocks[0].load().schema
        return Schema(self._schema) if self._schema is not None else None

    def columns(self) -> List[str]:
        s = self.schema()
        return s.names if s else []

    def get_internal_block_refs(self) -> List[Any]:
        return [b.ref for b in self._blocks]

    def to_arrow_refs(self) -> List[Any]:
        if all(b.offset == 0 and b.num_rows == b.stored_rows for b in self._blocks):
            return [b.ref for b in self._blocks]
        return [runtime.put(b.load().combine_chunks()) for b in self._blocks]

    def iter_internal_ref_bundles(self) -> Iterator[RefBundle]:
        for b in self._blocks:
            yield RefBundle([(b.ref, BlockMetadata(b.num_rows, b.size_bytes, self._schema))])

    def materialize(self) -> "Dataset":
        return self

    fully_executed = materialize

    def stats(self) -> str:
        return f"Dataset(num_blocks={self.num_blocks()}, num_rows={self.count()}, bytes={self.size_bytes()})"

    def __repr__(self):
        names = self.columns()
        return f"Dataset(num_blocks={self.num_blocks()}, num_rows={self.count()}, schema={names})"

    # ------------------------------------------------------------------ reading
    def iter_tables(self, pin: bool = False) -> Iterator[pa.Table]:
        for b in self._blocks:
            yield b.load(pin=pin)

    def to_arrow(self) -> pa.Table:
        tables = list(self.iter_tables())
        if not tables:
            return self._schema.empty_table() if self._schema is not None else pa.table({})
        return pa.concat_tables(tables)

    def to_pandas(self, limit: Optional[int] = None):
        table = self.to_arrow()
        return (table.slice(0, limit) if limit else table).to_pandas()

    def take(self, limit: int = 20) -> List[Dict[str, Any]]:
        out: List[Dict[str, Any]] = []
        for table in self.iter_tables():
            out.extend(table.slice(0, limit - len(out)).to_pylist())
            if len(out) >= limit:
                break
        return out

    def take_all(self) -> List[Dict[str, Any]]:
        return self.to_arrow().to_pylist()

    def show(self, limit: int = 20) -> None:
        for row in self.take(limit):
            print(row)

    def iter_rows(self) -> Iterator[Dict[str, Any]]:
        for table in self.iter_tables():
            yield from table.to_pylist()

    def iter_batches(self, batch_size: Optional[int] = 256, batch_format: str = "pyarrow",
                     drop_last: bool = False, prefetch_batches: int = 0,
                     local_shuffle_buffer_size: Optional[int] = None,
                     local_shuffle_seed: Optional[int] = None, **_ignored) -> Iterator[Any]:
        def column_numpy(col):
            arr = col.combine_chunks() if isinstance(col, pa.ChunkedArray) else col
            if pa.types.is_list(arr.type) or pa.types.is_large_list(arr.type) or pa.types.is_fixed_size_list(arr.type):
                rows = arr.to_pylist()                      # vectors: [n, width] when the rows share a width
                try:
                    return np.asarray(rows)
                except ValueError:
                    return np.asarray(rows, dtype=object)
            return arr.to_numpy(zero_copy_only=False)

        def convert(table: pa.Table):
            if batch_format in ("pyarrow", "arrow", "default"):
                return table
            if batch_format == "pandas":
                return table.to_pandas()
            if batch_format == "numpy":
                return {n: column_numpy(table.column(n)) for n in table.column_names}
            raise ValueError(f"unknown batch_format {batch_format!r}")
        if local_shuffle_buffer_size:
            # rows are drawn at random from a buffer of at least this many rows (plus the batch being cut)
            rng = np.random.default_rng(local_shuffle_seed)
            size = batch_size or 256
            pool: Optional[pa.Table] = None
            for table in self.iter_tables():
                pool = table if pool is None else pa.concat_tables([pool, table])
                while pool.num_rows >= local_shuffle_buffer_size + size:
                    pick = rng.choice(pool.num_rows, size, replace=False)
                    mask = np.ones(pool.num_rows, dtype=bool)
                    mask[pick] = False
                    yield convert(pool.take(pa.array(pick)))
                    pool = pool.filter(pa.array(mask))
            if pool is not None and pool.num_rows:
                pool = pool.take(pa.array(rng.permutation(pool.num_rows)))
                for off in builtins_range(0, pool.num_rows, size):
                    piece = pool.slice(off, size)
                    if piece.num_rows == size or not drop_last:
                        yield convert(piece)
            return
        if batch_size is None:
            for table in self.iter_tables():
                yield convert(table)
            return
        carry: List[pa.Table] = []
        have = 0
        for table in self.iter_tables():
            off = 0
            while off < table.num_rows:
                take = min(batch_size - have, table.num_rows - off)
                carry.append(table.slice(off, take))
                have += take
                off += take
                if have == batch_size:
                    yield convert(carry[0] if len(carry) == 1 else pa.concat_tables(carry))
                    carry, have = [], 0
        if have and not drop_last:
            yield convert(carry[0] if len(carry) == 1 else pa.concat_tables(carry))

    def iter_torch_batches(self, batch_size: int = 256, dtypes=None, device: Optional[str] = None,
                           drop_last: bool = False, collate_fn: Optional[Callable] = None,
                           local_shuffle_buffer_size: Optional[int] = None,
                           local_shuffle_seed: Optional[int] = None, prefetch_batches: int = 0,
                           **_ignored) -> Iterator[Dict[str, Any]]:
        import torch
        if collate_fn is None and not local_shuffle_buffer_size and batch_size:
            # blocks become tensors once; batches are slices (a batch spanning two blocks is stitched) -- per-batch
            # Arrow -> numpy conversion of every column dominates the loop at small batch sizes
            def block_tensors(table):
                out = {}
                for name in table.column_names:
                    col = table.column(name)
                    arr = col.combine_chunks() if isinstance(col, pa.ChunkedArray) else col
                    if pa.types.is_list(arr.type) or pa.types.is_large_list(arr.type) or pa.types.is_fixed_size_list(arr.type):
                        v = np.asarray(arr.to_pylist())
                    else:
                        v = arr.to_numpy(zero_copy_only=False)
                    if v.dtype == object:
                        return None                      # strings / ragged vectors: the generic path handles them
                    t = torch.as_tensor(np.ascontiguousarray(v))
                    if dtypes is not None:
                        t = t.to(dtypes[name] if isinstance(dtypes, dict) else dtypes)
                    out[name] = t
                return out

            def emit(d):
                return {k: (v.to(device) if device else v) for k, v in d.items()}
            carry, fallback, started = None, False, False
            for table in self.iter_tables():
                if table.num_rows == 0:
                    continue
                cur = block_tensors(table)
                if cur is None:
                    fallback = True
                    break
                started = True
                n, off = table.num_rows, 0
                if carry is not None:
                    have = next(iter(carry.values())).shape[0]
                    take = min(batch_size - have, n)
                    carry = {k: torch.cat([carry[k], cur[k][:take]]) for k in cur}
                    off = take
                    if have + take < batch_size:
                        continue
                    yield emit(carry)
                    carry = None
                while n - off >= batch_size:
                    yield emit({k: v[off:off + batch_size] for k, v in cur.items()})
                    off += batch_size
                if off < n:
                    carry = {k: v[off:] for k, v in cur.items()}
            if not fallback:
                if carry is not None and not drop_last:
                    yield emit(carry)
                return
            if started:                                  # cannot resume mid-stream on the generic path
                raise TypeError("iter_torch_batches: a later block holds non-numeric columns; pass collate_fn")
        for batch in self.iter_batches(batch_size, "numpy", drop_last,
                                       local_shuffle_buffer_size=local_shuffle_buffer_size,
                                       local_shuffle_seed=local_shuffle_seed):
            if collate_fn is not None:
                yield collate_fn(batch)
                continue
            out = {}
            for k, v in batch.items():
                t = torch.as_tensor(np.ascontiguousarray(v))
                if dtypes is not None:
                    t = t.to(dtypes[k] if isinstance(dtypes, dict) else dtypes)
                out[k] = t.to(device) if device else t
            yield out

    def to_torch(self, label_column: Optional[str] = None,
                 feature_columns: Optional[Union[List[str], List[List[str]], Dict[str, List[str]]]] = None,
                 label_column_dtype=None, feature_column_dtypes=None, batch_size: int = 1,
                 prefetch_batches: int = 1, drop_last: bool = False, unsqueeze_label_tensor: bool = True,
                 unsqueeze_feature_tensors: bool = True, null_fill: float = float("nan")):
        """Iterable of ``(features, label)`` torch tensors.

        Single dtype (or None) -> features is one ``[B, F]`` tensor; a list of per-column
        dtypes -> a list of ``[B, 1]`` tensors.  A NULL in a column that
        ends up as a floating tensor becomes ``null_fill`` -- NaN by default, which is what the
        reference's Arrow -> pandas -> numpy path yields; integer tensors get 0.  The GPU loader
        (``raydp_b200.ops.ingest``) follows the same rule inside its kernel.
        """
        from raydp_b200.data.torch_iter import TorchBatchIterable
        return TorchBatchIterable(self, label_column, feature_columns, label_column_dtype,
                                  feature_column_dtypes, batch_size, drop_last,
                                  unsqueeze_label_tensor, null_fill=null_fill)

    # ------------------------------------------------------------------ transforms
    def split(self, n: int, equal: bool = False, locality_hints=None) -> List["Dataset"]:
        if n <= 0:
            raise ValueError("n must be positive")
        if not equal:
            if self.num_blocks() >= n:
                per, rem = divmod(self.num_blocks(), n)
                out, off = [], 0
                for i in builtins_range(n):
                    size = per + (1 if i < rem else 0)
                    out.append(Dataset(self._blocks[off:off + size], self._schema))
                    off += size
                return out
            total = self.count()
            cuts = [int(round(total * (i + 1) / n)) for i in builtins_range(n - 1)]
            return self.split_at_indices(cuts)
        total = self.count()
        per = total // n
        pieces = self.split_at_indices([per * (i + 1) for i in builtins_range(n)])
        return pieces[:n]           # the remainder (< n rows) is dropped so ranks stay in step

    def split_at_indices(self, indices: List[int]) -> List["Dataset"]:
        if any(i < 0 for i in indices) or list(indices) != sorted(indices):
            raise ValueError("indices must be non-negative and sorted")
        bounds = [0] + list(indices) + [self.count()]
        out: List[Dataset] = []
        blk, off = 0, 0
        for lo, hi in zip(bounds[:-1], bounds[1:]):
            need = max(0, min(hi, self.count()) - lo)
            views: List[Block] = []
            while need > 0 and blk < len(self._blocks):
                b = self._blocks[blk]
                avail = b.num_rows - off
                if avail <= 0:
                    blk, off = blk + 1, 0
                    continue
                take = min(avail, need)
                frac = take / max(1, b.num_rows)
                views.append(Block(b.ref, b.offset + off, take, b.stored_rows,
                                   int(b.size_bytes * frac)))
                off += take
                need -= take
            out.append(Dataset(views, self._schema))
        return out

    def split_proportionately(self, proportions: List[float]) -> List["Dataset"]:
        total = self.count()
        acc, cuts = 0.0, []
        for p in proportions:
            acc += p
            cuts.append(int(total * acc))
        return self.split_at_indices(cuts)

    def train_test_split(self, test_size: float, shuffle: bool = False, seed: Optional[int] = None):
        ds = self.random_shuffle(seed) if shuffle else self
        a, b = ds.split_proportionately([1.0 - test_size])
        return a, b

    def random_shuffle(self, seed: Optional[int] = None, num_blocks: Optional[int] = None) -> "Dataset":
        """Global shuffle: map (random partition of every block) then reduce (concat + permute)."""
        if not self._blocks:
            return self
        seed = int(seed) if seed is not None else int(np.random.SeedSequence().entropy % (2 ** 31))
        n_out = num_blocks or self.num_blocks()
        blocks = self._blocks
        maps = _in_waves(lambda i: _shuffle_map.remote(blocks[i].ref, blocks[i].offset,
                                                       blocks[i].num_rows, n_out, seed, i),
                         len(blocks))
        pieces = runtime.get(maps)
        reduces = _in_waves(lambda d: _shuffle_reduce.remote(seed, d, *[row[d] for row in pieces]), n_out)
        intermediates = [p for row in pieces for p in row if p is not None]
        out = from_arrow_refs([r for r in reduces], _drop_empty=True, _schema=self._schema)
        runtime.free(intermediates)
        return out

    def randomize_block_order(self, seed: Optional[int] = None) -> "Dataset":
        order = np.random.default_rng(seed).permutation(len(self._blocks))
        return Dataset([self._blocks[i] for i in order], self._schema)

    def repartition(self, num_blocks: int) -> "Dataset":
        total = self.count()
        cuts = [int(round(total * (i + 1) / num_blocks)) for i in builtins_range(num_blocks - 1)]
        parts = self.split_at_indices(cuts)
        refs = [runtime.put(p.to_arrow().combine_chunks()) for p in parts]
        return from_arrow_refs(refs, _schema=self._schema)

    def map_batches(self, fn: Callable, batch_format: str = "pyarrow", **_kw) -> "Dataset":
        blocks = self._blocks
        refs = _in_waves(lambda i: _map_block.remote(fn, blocks[i].ref, blocks[i].offset,
                                                     blocks[i].num_rows, batch_format), len(blocks))
        return from_arrow_refs(refs)

    def map(self, fn: Callable[[Dict[str, Any]], Dict[str, Any]]) -> "Dataset":
        """Row-wise ``fn(row_dict) -> row_dict`` (``ray.data.Dataset.map``)."""
        def per_block(table: pa.Table, _fn=fn) -> pa.Table:
            rows = [_fn(r) for r in table.to_pylist()]
            return pa.Table.from_pylist(rows) if rows else table.slice(0, 0)
        return self.map_batches(per_block)

    def filter(self, fn: Callable[[Dict[str, Any]], bool]) -> "Dataset":
        """Keep the rows for which ``fn(row_dict)`` is true."""
        def per_block(table: pa.Table, _fn=fn) -> pa.Table:
            keep = pa.array([bool(_fn(r)) for r in table.to_pylist()], type=pa.bool_())
            return table.filter(keep)
        return self.map_batches(per_block)

    def add_column(self, name: str, fn: Callable) -> "Dataset":
        """``fn(pandas.DataFrame) -> Series`` evaluated per block and appended as column ``name``."""
        def per_block(table: pa.Table, _fn=fn, _name=name) -> pa.Table:
            return table.append_column(_name, pa.array(_fn(table.to_pandas())))
        return self.map_batches(per_block)

    def sort(self, key: str, descending: bool = False) -> "Dataset":
        """Global sort by one column (gathers on the driver; meant for small result sets)."""
        table = self.to_arrow().sort_by([(key, "descending" if descending else "ascending")])
        return from_arrow([table])

    def select_columns(self, cols: List[str]) -> "Dataset":
        return self.map_batches(lambda t, _c=tuple(cols): t.select(list(_c)))

    def drop_columns(self, cols: List[str]) -> "Dataset":
        return self.map_batches(lambda t, _c=tuple(cols): t.drop_columns(list(_c)))

    def limit(self, n: int) -> "Dataset":
        return self.split_at_indices([n])[0]

    def union(self, *others: "Dataset") -> "Dataset":
        blocks = list(self._blocks)
        for o in others:
            blocks.extend(o._blocks)
        return Dataset(blocks, self._schema)

    # ------------------------------------------------------------------ output
    def to_spark(self, spark, parallelism=None):
        from raydp_b200.spark.dataset import ray_dataset_to_spark_dataframe
        return ray_dataset_to_spark_dataframe(spark, self.schema(), self.to_arrow_refs())

    def to_tf(self, feature_columns, label_columns, batch_size: int = 1, drop_last: bool = False,
              feature_type_spec=None, label_type_spec=None):
        """``tf.data.Dataset`` of ``(features, labels)`` batches (``ray.data.Dataset.to_tf`` as consumed by
        reference ``tf/estimator.py:171-210``).  A single column name gives one tensor, a list gives a
        dict of per-column tensors.  Needs TensorFlow."""
        try:
            import tensorflow as tf
        except ImportError as exc:
            raise ImportError("Dataset.to_tf needs TensorFlow, which is not installed") from exc
        table = self.to_arrow()

        def pick(cols):
            if isinstance(cols, str):
                return table.column(cols).to_numpy(zero_copy_only=False)
            return {c: table.column(c).to_numpy(zero_copy_only=False) for c in cols}
        ds = tf.data.Dataset.from_tensor_slices((pick(feature_columns), pick(label_columns)))
        return ds.batch(batch_size, drop_remainder=drop_last)

    def write_parquet(self, path: str) -> None:
        import pyarrow.parquet as pq
        os.makedirs(path, exist_ok=True)
        for i, table in enumerate(self.iter_tables()):
            pq.write_table(table, os.path.join(path, f"block-{i:05d}.parquet"))




# --------------------------------------------------------------------------- constructors
def from_arrow_refs(refs: Sequence[Any], metadata: Optional[List[BlockMetadata]] = None,
                    _drop_empty: bool = False, _schema: Optional[pa.Schema] = None) -> Dataset:
    """Dataset over existing ``ObjectRef[pa.Table]`` blocks (no copy)."""
    refs = list(refs)
    if metadata is None:
        stats = runtime.get(_in_waves(lambda i: _stat_block.remote(refs[i]), len(refs))) if refs else []
        metadata = []
        for st in stats:
            if st is None:
                metadata.append(BlockMetadata(0, 0))
            else:
                metadata.append(BlockMetadata(st[0], st[1], pa.ipc.read_schema(pa.py_buffer(st[2]))))
    blocks, schema = [], _schema
    for ref, md in zip(refs, metadata):
        if _drop_empty and md.num_rows == 0:
            continue
        blocks.append(Block(ref, 0, md.num_rows, md.num_rows, md.size_bytes))
        if schema is None and md.schema is not None:
            schema = md.schema
    return Dataset(blocks, schema)


def from_arrow(tables: Union[pa.Table, List[pa.Table]]) -> Dataset:
    tables = [tables] if isinstance(tables, pa.Table) else list(tables)
    refs = [runtime.put(t) for t in tables]
    md = [BlockMetadata(t.num_rows, int(t.nbytes), t.schema) for t in tables]
    return from_arrow_refs(refs, md)


def from_pandas(dfs) -> Dataset:
    import pandas as pd
    dfs = [dfs] if isinstance(dfs, pd.DataFrame) else list(dfs)
    return from_arrow([pa.Table.from_pandas(df, preserve_index=False) for df in dfs])


def from_numpy(arr: np.ndarray, column: str = "data") -> Dataset:
    return from_arrow(pa.table({column: list(arr) if arr.ndim > 1 else arr}))


def from_items(items: List[Any], parallelism: int = 1) -> Dataset:
    rows = [i if isinstance(i, dict) else {"item": i} for i in items]
    n = max(1, min(parallelism, max(1, len(rows))))
    per = math.ceil(len(rows) / n) if rows else 0
    tables = [pa.Table.from_pylist(rows[i * per:(i + 1) * per]) for i in builtins_range(n)] if rows \
        else [pa.table({"item": pa.array([], pa.int64())})]
    return from_arrow([t for t in tables if t.num_rows or len(tables) == 1])


def range(n: int, parallelism: int = 4) -> Dataset:  # noqa: A001
    parallelism = max(1, min(parallelism, max(1, n)))
    bounds = np.linspace(0, n, parallelism + 1).astype(np.int64)
    return from_arrow([pa.table({"id": np.arange(bounds[i], bounds[i + 1], dtype=np.int64)})
                       for i in builtins_range(parallelism)])


def from_spark(df, parallelism: Optional[int] = None) -> Dataset:
    from raydp_b200.spark.dataset import spark_dataframe_to_ray_dataset
    return spark_dataframe_to_ray_dataset(df, parallelism)


def read_parquet(path: Union[str, List[str]], columns: Optional[List[str]] = None,
                 file_extensions: Optional[List[str]] = None, **_kw) -> Dataset:
    from raydp_b200.sql.plan import list_data_files
    paths = [path] if isinstance(path, str) else list(path)
    suffixes = tuple("." + e.lstrip(".") for e in (file_extensions or ["parquet"]))
    files = [f for p in paths for f in list_data_files(p, suffixes)]
    refs = _in_waves(lambda i: _read_parquet_file.remote(files[i], columns), len(files))
    return from_arrow_refs(refs)


def read_csv(path: Union[str, List[str]], **_kw) -> Dataset:
    from raydp_b200.sql.plan import list_data_files
    paths = [path] if isinstance(path, str) else list(path)
    files = [f for p in paths for f in list_data_files(p, ())]
    refs = _in_waves(lambda i: _read_csv_file.remote(files[i]), len(files))
    return from_arrow_refs(refs)


def read_json(path: Union[str, List[str]], **_kw) -> Dataset:
    """JSON-lines files, one block per file."""
    import pyarrow.json as pajson
    from raydp_b200.sql.plan import list_data_files
    paths = [path] if isinstance(path, str) else list(path)
    files = [f for p in paths for f in list_data_files(p, ())]
    ds = from_arrow([pajson.read_json(f) for f in files])
    ds._input_files = files
    return ds


def read_text(path: Union[str, List[str]], encoding: str = "utf-8", drop_empty_lines: bool = True, **_kw) -> Dataset:
    from raydp_b200.sql.plan import list_data_files
    paths = [path] if isinstance(path, str) else list(path)
    files = [f for p in paths for f in list_data_files(p, ())]
    tables = []
    for f in files:
        with open(f, "r", encoding=encoding) as fh:
            lines = fh.read().splitlines()
        tables.append(pa.table({"text": pa.array([ln for ln in lines if ln or not drop_empty_lines], pa.string())}))
    ds = from_arrow(tables)
    ds._input_files = files
    return ds


def read_numpy(path: Union[str, List[str]], **_kw) -> Dataset:
    from raydp_b200.sql.plan import list_data_files
    paths = [path] if isinstance(path, str) else list(path)
    files = [f for p in paths for f in list_data_files(p, (".npy",))]
    tables = []
    for f in files:
        arr = np.load(f)
        tables.append(pa.table({"data": list(arr) if arr.ndim > 1 else arr}))
    ds = from_arrow(tables)
    ds._input_files = files
    return ds


def read_binary_files(path: Union[str, List[str]], include_paths: bool = False, **_kw) -> Dataset:
    from raydp_b200.sql.plan import list_data_files
    paths = [path] if isinstance(path, str) else list(path)
    files = [f for p in paths for f in list_data_files(p, ())]
    tables = []
    for f in files:
        with open(f, "rb") as fh:
            cols = {"bytes": pa.array([fh.read()], pa.binary())}
        if include_paths:
            cols["path"] = pa.array([f])
        tables.append(pa.table(cols))
    ds = from_arrow(tables)
    ds._input_files = files
    return ds


def read_images(path: Union[str, List[str]], **_kw) -> Dataset:
    try:
        from PIL import Image
    except ImportError as exc:
        raise ImportError("read_images needs Pillow, which is not installed") from exc
    from raydp_b200.sql.plan import list_data_files
    paths = [path] if isinstance(path, str) else list(path)
    files = [f for p in paths for f in list_data_files(p, ())]
    return from_items([{"image": np.asarray(Image.open(f)).tolist(), "path": f} for f in files])


def range_tensor(n: int, shape=(1,), parallelism: int = 4) -> Dataset:
    shape = tuple(shape)
    parallelism = max(1, min(parallelism, max(1, n)))
    bounds = np.linspace(0, n, parallelism + 1).astype(np.int64)
    tables = []
    for i in builtins_range(parallelism):
        ids = np.arange(bounds[i], bounds[i + 1], dtype=np.int64)
        data = np.broadcast_to(ids.reshape((-1,) + (1,) * len(shape)), (len(ids),) + shape)
        tables.append(pa.table({"data": [row.tolist() for row in data]}))
    return from_arrow(tables)


def from_torch(dataset) -> Dataset:
    """A map-style ``torch.utils.data.Dataset`` -> one ``item`` column."""
    return from_items([_plain_item(dataset[i]) for i in builtins_range(len(dataset))])


def _plain_item(value):
    if hasattr(value, "tolist"):
        return value.tolist()
    if isinstance(value, (tuple, list)):
        return [_plain_item(v) for v in value]
    if isinstance(value, dict):
        return {k: _plain_item(v) for k, v in value.items()}
    return value


def from_huggingface(dataset) -> Dataset:
    """A ``datasets.Dataset`` (Arrow-backed) -> Dataset, without copying through Python objects."""
    table = getattr(getattr(dataset, "data", None), "table", None)
    if isinstance(table, pa.Table):
        return from_arrow([table.combine_chunks()])
    return from_pandas(dataset.to_pandas())


def from_pandas_refs(refs) -> Dataset:
    refs = [refs] if not isinstance(refs, (list, tuple)) else list(refs)
    return from_pandas([runtime.get(r) for r in refs])


def from_numpy_refs(refs, column: str = "data") -> Dataset:
    refs = [refs] if not isinstance(refs, (list, tuple)) else list(refs)
    return from_arrow([pa.table({column: list(a) if a.ndim > 1 else a}) for a in (runtime.get(r) for r in refs)])


def from_blocks(blocks: List[pa.Table]) -> Dataset:
    return from_arrow(list(blocks))


from raydp_b200.data import extras as _extras  # noqa: E402
from raydp_b200.data.extras import (AbsMax, AggregateFn, Count, DataContext, DataIterator, GroupedData, Max,  # noqa: E402,F401
                                    Mean, Min, Quantile, Std, Sum, Unique)

_extras._install(Dataset)
