"""Physical plan of the built-in DataFrame engine: a partitioned source + narrow operators.

This is the part of the reference's JVM layer that matters for training: turning rows into
Arrow record batches on executors (reference ``ObjectStoreWriter.scala:172-174``
``df.toArrowBatchRdd``).  Here partitions *are* Arrow tables from the start, every narrow
operator is a ``pyarrow.compute`` call on one partition, and wide operators (repartition,
groupBy, join, sort) are two-stage jobs assembled by ``dataframe.py``.
"""
from __future__ import annotations

import glob as _glob
import os
import re
from typing import Any, Callable, Dict, List, Optional, Sequence, Tuple

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

from raydp_b200.sql import types as T
from raydp_b200.sql.column import Column, EvalContext, _full, parse_timestamp_strings


def strip_scheme(path: str) -> str:
    return re.sub(r"^file:(//)?", "", path)


# --------------------------------------------------------------------------- sources
class Source:
    num_partitions: int = 1

    def load(self, idx: int) -> pa.Table:
        raise NotImplementedError

    def schema(self) -> pa.Schema:
        return self.load(0).schema


class RangeSource(Source):
    def __init__(self, start: int, end: int, step: int, num_partitions: int):
        self.start, self.end, self.step = int(start), int(end), int(step)
        total = max(0, -(-(self.end - self.start) // self.step))
        self.total = total
        self.num_partitions = max(1, min(int(num_partitions), max(1, total)))

    def bounds(self, idx: int) -> Tuple[int, int]:
        per, rem = divmod(self.total, self.num_partitions)
        lo = idx * per + min(idx, rem)
        return lo, lo + per + (1 if idx < rem else 0)

    def load(self, idx: int) -> pa.Table:
        lo, hi = self.bounds(idx)
        ids = np.arange(self.start + lo * self.step, self.start + hi * self.step, self.step,
                        dtype=np.int64)
        return pa.table({"id": ids})

    def schema(self) -> pa.Schema:
        return pa.schema([pa.field("id", pa.int64(), nullable=False)])


class LocalSource(Source):
    """Tables that travel inside the plan (``createDataFrame`` of driver-side data)."""

    def __init__(self, tables: List[pa.Table]):
        self.tables = tables
        self.num_partitions = max(1, len(tables))

    def load(self, idx: int) -> pa.Table:
        return self.tables[idx]

    def schema(self) -> pa.Schema:
        return self.tables[0].schema


class RefSource(Source):
    """Partitions stored in the object store: partition i = concat(refs[i])."""

    def __init__(self, refs: List[List[Any]], schema: pa.Schema):
        self.refs = refs
        self._schema = schema
        self.num_partitions = max(1, len(refs))

    def load(self, idx: int) -> pa.Table:
        from raydp_b200 import runtime
        if not self.refs:
            return self._schema.empty_table()
        parts = [runtime.get(r) for r in self.refs[idx]]
        parts = [p if isinstance(p, pa.Table) else pa.ipc.open_stream(p).read_all() for p in parts]
        if not parts:
            return self._schema.empty_table()
        return parts[0] if len(parts) == 1 else pa.concat_tables(parts)

    def schema(self) -> pa.Schema:
        return self._schema


_TS_RE = re.compile(r"^\d{4}-\d{2}-\d{2}[ T]\d{2}:\d{2}:\d{2}(\.\d+)?( ?(UTC|Z))?$")


def _infer_timestamps(table: pa.Table) -> pa.Table:
    for i, field in enumerate(table.schema):
        if not pa.types.is_string(field.type) or table.num_rows == 0:
            continue
        sample = [v for v in table.column(i).slice(0, 32).to_pylist() if v is not None]
        if sample and all(_TS_RE.match(v) and "." not in v for v in sample):
            table = table.set_column(i, field.name, parse_timestamp_strings(table.column(i)))
    return table


class CsvSource(Source):
    def __init__(self, files: List[str], header: bool, infer: bool, sep: str,
                 schema: Optional[T.StructType]):
        self.files = files
        self.header, self.infer, self.sep = header, infer, sep
        self.user_schema = schema
        self.num_partitions = max(1, len(files))
        self._cached_schema: Optional[pa.Schema] = None

    def load(self, idx: int) -> pa.Table:
        import pyarrow.csv as pacsv
        names = None if self.header else (
            self.user_schema.names if self.user_schema is not None else None)
        read_opts = pacsv.ReadOptions(column_names=names, autogenerate_column_names=(
            names is None and not self.header))
        parse_opts = pacsv.ParseOptions(delimiter=self.sep)
        types = None
        if self.user_schema is not None:
            types = {f.name: f.dataType.arrow for f in self.user_schema}
        elif not self.infer:
            # Spark without inferSchema reads everything as string
            head = pacsv.read_csv(self.files[idx], read_options=read_opts, parse_options=parse_opts)
            types = {n: pa.string() for n in head.column_names}
        conv = pacsv.ConvertOptions(column_types=types, strings_can_be_null=True,
                                    null_values=["", "NULL", "null", "NaN"] if types is None else [""])
        table = pacsv.read_csv(self.files[idx], read_options=read_opts, parse_options=parse_opts,
                               convert_options=conv)
        if self.user_schema is None and self.infer:
            table = _infer_timestamps(table)
            cols = []
            for f in table.schema:  # Spark infers int (32-bit) when values fit
                c = table.column(f.name)
                if pa.types.is_int64(f.type) and table.num_rows:
                    mm = pc.min_max(c).as_py()
                    if mm["min"] is not None and -2 ** 31 <= mm["min"] and mm["max"] < 2 ** 31:
                        c = pc.cast(c, pa.int32())
                cols.append(c)
            table = pa.table(cols, names=table.column_names)
        if not self.header and names is None:
            table = table.rename_columns([f"_c{i}" for i in range(table.num_columns)])
        return table

    def schema(self) -> pa.Schema:
        if self._cached_schema is None:
            self._cached_schema = self.load(0).schema
        return self._cached_schema


class ParquetSource(Source):
    def __init__(self, files: List[str], columns: Optional[List[str]] = None):
        self.files = files
        self.columns = columns
        self.num_partitions = max(1, len(files))

    def load(self, idx: int) -> pa.Table:
        import pyarrow.parquet as pq
        # the file alone: read_table() would also turn key=value directory names into columns
        return pq.ParquetFile(self.files[idx]).read(columns=self.columns)

    def schema(self) -> pa.Schema:
        import pyarrow.parquet as pq
        return pq.read_schema(self.files[0])


def partition_values(path: str, root: Optional[str] = None) -> List[Tuple[str, str]]:
    """Hive-style partition directories in a file path: ``.../year=2020/month=1/part-0.parquet``."""
    rel = os.path.relpath(path, root) if root else path
    out = []
    for part in os.path.dirname(rel).split(os.sep):
        if "=" in part and not part.startswith(("_", ".")):
            k, _, v = part.partition("=")
            out.append((k, v))
    return out


def _typed_partition_column(values: List[Optional[str]]) -> pa.Array:
    """Partition values are text in the path; ints and floats are recognised like Spark's inference does."""
    live = [v for v in values if v is not None]
    for caster, arrow_type in ((int, pa.int32()), (float, pa.float64())):
        try:
            converted = [None if v is None else caster(v) for v in values]
            if arrow_type == pa.int32() and any(abs(x) >= 2 ** 31 for x in converted if x is not None):
                arrow_type = pa.int64()
            if live:
                return pa.array(converted, type=arrow_type)
        except ValueError:
            continue
    return pa.array(values, type=pa.string())


class FileSource(Source):
    """One partition per file for the line / columnar formats: ``parquet``, ``orc``, ``json`` (one object
    per line), ``text``.  Columns encoded in the directory names (``key=value``) are appended."""

    def __init__(self, files: List[str], fmt: str, root: Optional[str] = None,
                 schema: Optional[T.StructType] = None, options: Optional[Dict[str, str]] = None):
        self.files, self.fmt, self.root = files, fmt, root
        self.user_schema = schema
        self.options = dict(options or {})
        self.num_partitions = max(1, len(files))
        self._cached_schema: Optional[pa.Schema] = None
        keys: List[str] = []
        for f in files:
            for k, _ in partition_values(f, root):
                if k not in keys:
                    keys.append(k)
        self.partition_keys = keys
        self._partition_types: Dict[str, pa.DataType] = {}
        for k in keys:
            vals = [dict(partition_values(f, root)).get(k) for f in files]
            vals = [None if v == "__HIVE_DEFAULT_PARTITION__" else v for v in vals]
            self._partition_types[k] = _typed_partition_column(vals).type

    def _read(self, path: str) -> pa.Table:
        if self.fmt == "parquet":
            import pyarrow.parquet as pq
            return pq.ParquetFile(path).read()       # the file alone: partition columns are added below
        if self.fmt == "orc":
            import pyarrow.orc as orc
            return orc.read_table(path)
        if self.fmt == "json":
            import pyarrow.json as pajson
            explicit = None
            if self.user_schema is not None:
                explicit = pa.schema([(f.name, f.dataType.arrow) for f in self.user_schema
                                      if f.name not in self.partition_keys])
            if os.path.getsize(path) == 0:
                return (explicit or pa.schema([])).empty_table()
            table = pajson.read_json(path, parse_options=pajson.ParseOptions(
                explicit_schema=explicit, unexpected_field_behavior="ignore" if explicit is not None else "infer"))
            return _infer_timestamps(table) if explicit is None else table
        if self.fmt == "text":
            with open(path, "r", encoding="utf-8") as fh:
                if str(self.options.get("wholetext", "false")).lower() == "true":
                    lines = [fh.read()]
                else:
                    lines = fh.read().splitlines()
            return pa.table({"value": pa.array(lines, type=pa.string())})
        raise NotImplementedError(f"data source format {self.fmt!r} is not supported")

    def load(self, idx: int) -> pa.Table:
        if not self.files:
            return self.schema().empty_table()
        path = self.files[idx]
        table = self._read(path)
        if self.fmt == "json" and self.user_schema is None and len(self.files) > 1:
            # objects omit null fields, so files can disagree on the columns: align to the union
            if getattr(self, "_json_schema", None) is None:
                self._json_schema = pa.unify_schemas([self._read(f).schema for f in self.files],
                                                     promote_options="permissive")
            cols = [table.column(f.name).cast(f.type) if f.name in table.column_names
                    else pa.nulls(table.num_rows, f.type) for f in self._json_schema]
            table = pa.table(cols, names=self._json_schema.names)
        if self.user_schema is not None and self.fmt != "json":
            wanted = [f for f in self.user_schema if f.name not in self.partition_keys]
            table = table.select([f.name for f in wanted]).cast(
                pa.schema([(f.name, f.dataType.arrow) for f in wanted]), safe=False)
        found = dict(partition_values(path, self.root))
        for k in self.partition_keys:
            v = found.get(k)
            v = None if v == "__HIVE_DEFAULT_PARTITION__" else v
            t = self._partition_types[k]
            value = None if v is None else (int(v) if pa.types.is_integer(t) else float(v) if pa.types.is_floating(t) else v)
            column = pa.array([value] * table.num_rows, type=t)
            if k in table.column_names:
                table = table.set_column(table.column_names.index(k), k, column)
            else:
                table = table.append_column(k, column)
        return table

    def schema(self) -> pa.Schema:
        if self._cached_schema is None:
            if not self.files:
                self._cached_schema = self.user_schema.to_arrow_schema() if self.user_schema is not None else pa.schema([])
            else:
                self._cached_schema = self.load(0).schema
        return self._cached_schema


class FuncSource(Source):
    """``parallelize`` / RDD-style source: partition i = fn(i) -> pa.Table."""

    def __init__(self, fn: Callable[[int], pa.Table], num_partitions: int, schema: pa.Schema):
        self.fn = fn
        self.num_partitions = num_partitions
        self._schema = schema

    def load(self, idx: int) -> pa.Table:
        return self.fn(idx)

    def schema(self) -> pa.Schema:
        return self._schema


def list_data_files(path: str, suffixes: Sequence[str]) -> List[str]:
    path = strip_scheme(path)
    if os.path.isdir(path):
        files = sorted(f for f in _glob.glob(os.path.join(path, "**", "*"), recursive=True)
                       if os.path.isfile(f) and not os.path.basename(f).startswith(("_", "."))
                       and (not suffixes or f.endswith(tuple(suffixes))))
    else:
        files = sorted(_glob.glob(path)) or [path]
    if not files:
        raise FileNotFoundError(f"Path does not exist: {path}")
    return files


# --------------------------------------------------------------------------- operators
def _apply(op: tuple, table: pa.Table, ctx: EvalContext) -> pa.Table:
    kind = op[0]
    n = table.num_rows
    if kind == "with_column":
        _, name, expr = op
        arr = _full(expr.evaluate(table, ctx), n)
        if name in table.column_names:
            return table.set_column(table.column_names.index(name), name, arr)
        return table.append_column(name, arr)
    if kind == "filter":
        mask = _full(op[1].evaluate(table, ctx), n)
        return table.filter(pc.fill_null(mask, False))
    if kind == "select":
        arrays, names = [], []
        for expr in op[1]:
            if expr._kind == "col" and expr._args[0] == "*":
                arrays.extend(table.columns)
                names.extend(table.column_names)
                continue
            arrays.append(_full(expr.evaluate(table, ctx), n))
            names.append(expr._output_name())
        return pa.table(arrays, names=names) if arrays else table.select([])
    if kind == "drop":
        keep = [i for i, c in enumerate(table.column_names) if c not in op[1]]   # by index: names may repeat
        return table.select(keep)
    if kind == "rename":
        return table.rename_columns([op[2] if c == op[1] else c for c in table.column_names])
    if kind == "rename_all":
        return table.rename_columns(list(op[1]))
    if kind == "fillna":
        _, value, subset = op
        cols = []
        for f in table.schema:
            c = table.column(f.name)
            if (subset is None or f.name in subset) and c.null_count:
                v = value.get(f.name) if isinstance(value, dict) else value
                if v is not None and _compatible(v, f.type):
                    c = pc.fill_null(c, pa.scalar(v).cast(f.type))
            cols.append(c)
        return pa.table(cols, names=table.column_names)
    if kind == "dropna":
        _, subset = op
        mask = None
        for name in (subset or table.column_names):
            valid = pc.is_valid(table.column(name))
            mask = valid if mask is None else pc.and_(mask, valid)
        return table.filter(mask) if mask is not None else table
    if kind == "sort_within":
        from raydp_b200.sql.windowing import sort_indices
        return table.take(sort_indices(table, op[1], ctx))
    if kind == "split":  # randomSplit piece: uniform draw in [lo, hi)
        _, lo, hi, seed = op
        rng = np.random.default_rng([int(seed) & 0x7FFFFFFF, ctx.partition_index, 7])
        draw = rng.random(n)
        return table.filter(pa.array((draw >= lo) & (draw < hi)))
    if kind == "sample":
        _, fraction, seed = op
        rng = np.random.default_rng([int(seed) & 0x7FFFFFFF, ctx.partition_index, 11])
        return table.filter(pa.array(rng.random(n) < fraction))
    if kind == "limit_local":
        return table.slice(0, op[1])
    if kind == "map_arrow":
        return op[1](table)
    if kind == "map_pandas":
        _, fn, schema = op
        import pandas as pd
        frames = list(fn(iter([table.to_pandas()])))
        if not frames:
            return schema.empty_table()
        out = pa.Table.from_pandas(pd.concat(frames, ignore_index=True), preserve_index=False)
        return out.select(schema.names).cast(schema, safe=False) if schema is not None else out
    if kind == "join":
        _, right, keys, how, right_keys = op
        how = {"inner": "inner", "left": "left outer", "left_outer": "left outer",
               "leftouter": "left outer", "right": "right outer", "right_outer": "right outer",
               "outer": "full outer", "full": "full outer", "full_outer": "full outer",
               "left_semi": "left semi", "leftsemi": "left semi", "left_anti": "left anti",
               "leftanti": "left anti"}[how]
        out = _join_tables(table, right, list(keys), list(right_keys), how)
        if how in ("left semi", "left anti") or len(set(out.column_names)) != len(out.column_names):
            return out
        # Spark's layout: join keys, then the other left columns, then the other right columns
        rest_left = [c for c in table.column_names if c not in keys]
        rest_right = [c for c in right.column_names if c not in right_keys and c not in table.column_names]
        wanted = [c for c in list(keys) + rest_left + rest_right if c in out.column_names]
        return out.select(wanted) if wanted != out.column_names else out
    if kind == "distinct":
        # Spark's equality for DISTINCT: NaN equals NaN and -0.0 equals 0.0
        for i, f in enumerate(table.schema):
            if pa.types.is_floating(f.type) and table.num_rows:
                col = table.column(i)
                col = pc.if_else(pc.is_nan(col), pa.scalar(float("nan"), f.type), pc.add(col, pa.scalar(0.0, f.type)))
                table = table.set_column(i, f, col)
        try:
            return table.group_by(table.column_names, use_threads=False).aggregate([])
        except (pa.ArrowNotImplementedError, pa.ArrowInvalid):
            # nested columns (arrays, structs, maps) are not hashable by Arrow's group-by: compare by value
            seen, keep = set(), []
            cols = [c.to_pylist() for c in table.columns]
            for i, row in enumerate(zip(*cols)):
                key = repr(row)
                if key not in seen:
                    seen.add(key)
                    keep.append(i)
            return table.take(pa.array(keep, type=pa.int64()))
    if kind == "distinct_subset":          # keep the first row of every key combination
        if table.num_rows == 0:
            return table
        tagged = table.append_column("__row", pa.array(np.arange(table.num_rows, dtype=np.int64)))
        first = tagged.group_by(list(op[1]), use_threads=False).aggregate([("__row", "min")])
        rows = first.column("__row_min").combine_chunks()
        return table.take(rows.take(pc.sort_indices(rows)))     # original order within the partition
    if kind == "cast_schema":
        return table.cast(op[1], safe=False)
    if kind == "generate":
        # select(<exprs...>, explode|posexplode[_outer](<array or map>), <exprs...>)
        _, exprs, gi = op
        gen = exprs[gi]
        outer = len(gen._args) > 1 and bool(gen._args[1])
        arr = _full(gen._args[0].evaluate(table, ctx), n)
        if isinstance(arr, pa.ChunkedArray):
            arr = arr.combine_chunks()
        is_map = pa.types.is_map(arr.type)
        lengths = np.asarray(pc.fill_null(pc.list_value_length(arr), 0)).astype(np.int64) if not is_map else \
            np.array([0 if v is None else len(v) for v in arr.to_pylist()], dtype=np.int64)
        out_len = np.where(lengths == 0, 1, lengths) if outer else lengths
        parent = np.repeat(np.arange(n, dtype=np.int64), out_len)
        starts_out = np.concatenate([[0], np.cumsum(out_len)[:-1]]) if n else np.zeros(0, dtype=np.int64)
        starts_in = np.concatenate([[0], np.cumsum(lengths)[:-1]]) if n else np.zeros(0, dtype=np.int64)
        pos = np.arange(len(parent), dtype=np.int64) - starts_out[parent]
        hollow = lengths[parent] == 0                      # rows produced only because of *_outer
        src = pa.array(np.where(hollow, 0, starts_in[parent] + pos), mask=hollow)
        if is_map:
            flat_k, flat_v = arr.keys[arr.offsets[0].as_py():], arr.items[arr.offsets[0].as_py():]
            produced = [("key", flat_k.take(src) if len(flat_k) else pa.nulls(len(parent), flat_k.type)),
                        ("value", flat_v.take(src) if len(flat_v) else pa.nulls(len(parent), flat_v.type))]
        elif gen._kind == "inline":              # array<struct<...>> -> one column per struct field
            flat = pc.list_flatten(arr)
            taken = flat.take(src) if len(flat) else pa.nulls(len(parent), flat.type)
            produced = [(f.name, pc.struct_field(taken, f.name)) for f in flat.type]
        else:
            flat = pc.list_flatten(arr)
            produced = [("col", flat.take(src) if len(flat) else pa.nulls(len(parent), flat.type))]
        if gen._kind == "posexplode":
            produced.insert(0, ("pos", pa.array(pos.astype(np.int32), mask=hollow)))
        alias = gen._name
        if alias is not None:
            names = list(alias) if isinstance(alias, (tuple, list)) else [alias]
            if len(names) != len(produced):
                raise ValueError(f"the generator produces {len(produced)} columns but {len(names)} aliases were given")
            produced = [(nm, v) for nm, (_, v) in zip(names, produced)]
        cols, names = [], []
        parent_idx = pa.array(parent)
        for i, e in enumerate(exprs):
            if i == gi:
                for nm, v in produced:
                    names.append(nm)
                    cols.append(v)
                continue
            value = _full(e.evaluate(table, ctx), n)
            names.append(e._output_name())
            cols.append(value.take(parent_idx))
        return pa.table(cols, names=names)
    raise NotImplementedError(f"operator {kind!r}")


def _nested(t: pa.DataType) -> bool:
    return pa.types.is_list(t) or pa.types.is_large_list(t) or pa.types.is_struct(t) or pa.types.is_map(t) or \
        pa.types.is_fixed_size_list(t)


def _join_tables(left: pa.Table, right: pa.Table, keys: List[str], right_keys: List[str], how: str) -> pa.Table:
    """Hash join.  Arrow's join kernel carries only flat payload columns, so when a side has arrays / structs /
    maps the join runs on (keys, row number) and the payload is gathered by row number afterwards."""
    payload_nested = any(_nested(f.type) for f in left.schema if f.name not in keys) or \
        any(_nested(f.type) for f in right.schema if f.name not in right_keys)
    if not payload_nested:
        return left.join(right, keys=keys, right_keys=right_keys, join_type=how, use_threads=False)
    slim_l = left.select(keys).append_column("__li", pa.array(np.arange(left.num_rows, dtype=np.int64)))
    slim_r = right.select(right_keys).append_column("__ri", pa.array(np.arange(right.num_rows, dtype=np.int64)))
    joined = slim_l.join(slim_r, keys=keys, right_keys=right_keys, join_type=how, use_threads=False)
    cols, names = [], []

    def gather(source: pa.Table, index_name: str, skip: List[str]) -> None:
        if index_name not in joined.column_names:
            return
        idx = joined.column(index_name).combine_chunks()
        for f in source.schema:
            if f.name in skip:
                continue
            names.append(f.name)
            cols.append(source.column(f.name).combine_chunks().take(idx))       # null index -> null value
    for k in keys:
        names.append(k)
        cols.append(joined.column(k))
    if how in ("right semi", "right anti"):
        gather(right, "__ri", right_keys)
    else:
        gather(left, "__li", keys)
        if how not in ("left semi", "left anti"):
            gather(right, "__ri", right_keys)
    return pa.table(cols, names=names)


def _compatible(value, arrow_type: pa.DataType) -> bool:
    if isinstance(value, bool):
        return pa.types.is_boolean(arrow_type)
    if isinstance(value, (int, float)):
        return pa.types.is_integer(arrow_type) or pa.types.is_floating(arrow_type)
    if isinstance(value, str):
        return pa.types.is_string(arrow_type)
    return False


class Plan:
    def __init__(self, source: Source, ops: Tuple[tuple, ...] = ()):
        self.source = source
        self.ops = tuple(ops)

    @property
    def num_partitions(self) -> int:
        return self.source.num_partitions

    def with_op(self, op: tuple) -> "Plan":
        return Plan(self.source, self.ops + (op,))

    def compute(self, idx: int) -> pa.Table:
        ctx = EvalContext(partition_index=idx)
        table = self.source.load(idx)
        for op in self.ops:
            table = _apply(op, table, ctx)
        return table

    def schema(self) -> pa.Schema:
        ctx = EvalContext(partition_index=0)
        table = self.source.schema().empty_table()
        for op in self.ops:
            table = _apply(op, table, ctx)
        return table.schema
