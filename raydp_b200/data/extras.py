"""The rest of the ``ray.data`` surface users of the reference reach for.

The reference moves Spark DataFrames into ``ray.data`` datasets (``raydp.spark.dataset``, reference
``python/raydp/spark/dataset.py:150-260``) and hands them to trainers; user programs then call the usual
dataset helpers on the result.  The core (:mod:`raydp_b200.data`) has blocks, splits, shuffles, batch
iteration and the torch bridge; this module adds aggregation (global and grouped), ``zip`` / ``join``,
more readers and writers, and a few conveniences.  Blocks are Arrow tables in the object store;
per-block work goes through ``map_batches`` (remote tasks), cross-block work merges small partial
results on the caller.
"""
from __future__ import annotations

import builtins
import os
from typing import Any, Callable, Dict, Iterator, List, Optional, Sequence, Union

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc


# ------------------------------------------------------------------------------------------ aggregation
class AggregateFn:
    """``ray.data.aggregate.AggregateFn``: ``init(key) -> acc``, ``accumulate_row(acc, row)``,
    ``merge(a, b)``, ``finalize(acc)``."""

    def __init__(self, init: Callable, merge: Callable, accumulate_row: Optional[Callable] = None,
                 accumulate_block: Optional[Callable] = None, finalize: Optional[Callable] = None,
                 name: Optional[str] = None):
        self.init, self.merge = init, merge
        self.accumulate_row, self.accumulate_block = accumulate_row, accumulate_block
        self.finalize = finalize or (lambda a: a)
        self.name = name or "agg"


class _Simple(AggregateFn):
    """Built-in aggregate over one column, computed with Arrow on each block and merged."""

    arrow_fn = ""

    def __init__(self, on: Optional[str] = None, ignore_nulls: bool = True, alias_name: Optional[str] = None, **kw):
        self.on, self.ignore_nulls, self.kw = on, ignore_nulls, kw
        self.name = alias_name or f"{type(self).__name__.lower()}({on or ''})"

    def of_table(self, table: pa.Table):
        raise NotImplementedError

    def combine(self, parts: list):
        raise NotImplementedError


class Count(_Simple):
    def of_table(self, table):
        return table.num_rows if self.on is None else len(table.column(self.on)) - table.column(self.on).null_count

    def combine(self, parts):
        return int(builtins.sum(parts))


class Sum(_Simple):
    def of_table(self, table):
        return pc.sum(table.column(self.on)).as_py()

    def combine(self, parts):
        live = [p for p in parts if p is not None]
        return builtins.sum(live) if live else None


class Min(_Simple):
    def of_table(self, table):
        return pc.min(table.column(self.on)).as_py()

    def combine(self, parts):
        live = [p for p in parts if p is not None]
        return builtins.min(live) if live else None


class Max(_Simple):
    def of_table(self, table):
        return pc.max(table.column(self.on)).as_py()

    def combine(self, parts):
        live = [p for p in parts if p is not None]
        return builtins.max(live) if live else None


class Mean(_Simple):
    def of_table(self, table):
        c = table.column(self.on)
        return (pc.sum(c).as_py() or 0, len(c) - c.null_count)

    def combine(self, parts):
        total, n = builtins.sum(p[0] for p in parts), builtins.sum(p[1] for p in parts)
        return total / n if n else None


class Std(_Simple):
    def of_table(self, table):
        x = pc.cast(table.column(self.on).drop_null(), pa.float64()).to_numpy()
        return (float(x.sum()), float((x * x).sum()), len(x))

    def combine(self, parts):
        s, sq, n = (builtins.sum(p[i] for p in parts) for i in builtins.range(3))
        ddof = self.kw.get("ddof", 1)
        if n - ddof <= 0:
            return None
        return float(np.sqrt(builtins.max(sq - s * s / n, 0.0) / (n - ddof)))


class AbsMax(_Simple):
    def of_table(self, table):
        return pc.max(pc.abs(table.column(self.on))).as_py()

    def combine(self, parts):
        live = [p for p in parts if p is not None]
        return builtins.max(live) if live else None


class Unique(_Simple):
    def of_table(self, table):
        return pc.unique(table.column(self.on).combine_chunks()).to_pylist()

    def combine(self, parts):
        seen: dict = {}
        for p in parts:
            for v in p:
                seen.setdefault(v, None)
        return list(seen)


class Quantile(_Simple):
    def __init__(self, on=None, q: float = 0.5, ignore_nulls: bool = True, alias_name=None):
        super().__init__(on, ignore_nulls, alias_name)
        self.q = q

    def of_table(self, table):
        return table.column(self.on).drop_null().to_numpy()

    def combine(self, parts):
        allv = np.concatenate(parts) if parts else np.zeros(0)
        return float(np.quantile(allv, self.q)) if len(allv) else None


def _aggregate_tables(tables: Sequence[pa.Table], aggs: Sequence[AggregateFn]) -> Dict[str, Any]:
    out: Dict[str, Any] = {}
    for agg in aggs:
        if isinstance(agg, _Simple):
            out[agg.name] = agg.combine([agg.of_table(t) for t in tables if t.num_rows or agg.on is None])
            continue
        partials = []
        for t in tables:
            acc = agg.init(None)
            if agg.accumulate_block is not None:
                acc = agg.accumulate_block(acc, t)
            else:
                for row in t.to_pylist():
                    acc = agg.accumulate_row(acc, row)
            partials.append(acc)
        merged = agg.init(None)
        for p in partials:
            merged = agg.merge(merged, p)
        out[agg.name] = agg.finalize(merged)
    return out


class GroupedData:
    """``ds.groupby(key)``: ``count / sum / min / max / mean / std / aggregate / map_groups``."""

    def __init__(self, ds, key: Union[str, List[str], None]):
        self._ds = ds
        self._keys = [] if key is None else ([key] if isinstance(key, str) else list(key))

    def _arrow(self, specs: List[tuple], names: List[str]):
        from raydp_b200.data import from_arrow
        table = self._ds.to_arrow()
        if self._keys:
            grouped = table.group_by(self._keys, use_threads=False).aggregate(specs)
            grouped = grouped.sort_by([(k, "ascending") for k in self._keys])
        else:
            grouped = pa.TableGroupBy(table, [], use_threads=False).aggregate(specs)
        produced = [f"{c}_{fn}" if not isinstance(c, list) else fn for c, fn in specs]
        cols = [grouped.column(k) for k in self._keys] + [grouped.column(p) for p in produced]
        return from_arrow([pa.table(cols, names=self._keys + names)])

    def _columns(self, on) -> List[str]:
        if on is None:
            schema = self._ds.to_arrow().schema
            return [f.name for f in schema if f.name not in self._keys and
                    (pa.types.is_integer(f.type) or pa.types.is_floating(f.type))]
        return [on] if isinstance(on, str) else list(on)

    def count(self):
        return self._arrow([([], "count_all")], ["count()"])

    def sum(self, on=None, ignore_nulls: bool = True):
        cols = self._columns(on)
        return self._arrow([(c, "sum") for c in cols], [f"sum({c})" for c in cols])

    def min(self, on=None, ignore_nulls: bool = True):
        cols = self._columns(on)
        return self._arrow([(c, "min") for c in cols], [f"min({c})" for c in cols])

    def max(self, on=None, ignore_nulls: bool = True):
        cols = self._columns(on)
        return self._arrow([(c, "max") for c in cols], [f"max({c})" for c in cols])

    def mean(self, on=None, ignore_nulls: bool = True):
        cols = self._columns(on)
        return self._arrow([(c, "mean") for c in cols], [f"mean({c})" for c in cols])

    def std(self, on=None, ddof: int = 1, ignore_nulls: bool = True):
        cols = self._columns(on)
        from raydp_b200.data import from_arrow
        table = self._ds.to_arrow()
        grouped = table.group_by(self._keys, use_threads=False).aggregate(
            [(c, "stddev", pc.VarianceOptions(ddof=ddof)) for c in cols]).sort_by([(k, "ascending") for k in self._keys])
        out = [grouped.column(k) for k in self._keys] + [grouped.column(f"{c}_stddev") for c in cols]
        return from_arrow([pa.table(out, names=self._keys + [f"std({c})" for c in cols])])

    def aggregate(self, *aggs: AggregateFn):
        from raydp_b200.data import from_arrow
        table = self._ds.to_arrow()
        if not self._keys:
            return from_arrow([pa.Table.from_pylist([_aggregate_tables([table], aggs)])])
        keys_tbl = table.group_by(self._keys, use_threads=False).aggregate([]).sort_by(
            [(k, "ascending") for k in self._keys])
        rows = []
        for key_row in keys_tbl.to_pylist():
            mask = None
            for k, v in key_row.items():
                hit = pc.is_null(table.column(k)) if v is None else pc.fill_null(pc.equal(table.column(k), v), False)
                mask = hit if mask is None else pc.and_(mask, hit)
            rows.append({**key_row, **_aggregate_tables([table.filter(mask)], aggs)})
        return from_arrow([pa.Table.from_pylist(rows)])

    def map_groups(self, fn: Callable, batch_format: str = "pandas", **_kw):
        from raydp_b200.data import from_arrow
        table = self._ds.to_arrow()
        pieces = []
        keys_tbl = table.group_by(self._keys, use_threads=False).aggregate([]).sort_by(
            [(k, "ascending") for k in self._keys]) if self._keys else pa.table({"__all": [0]})
        for key_row in keys_tbl.to_pylist():
            group = table
            if self._keys:
                mask = None
                for k, v in key_row.items():
                    hit = pc.is_null(table.column(k)) if v is None else pc.fill_null(pc.equal(table.column(k), v), False)
                    mask = hit if mask is None else pc.and_(mask, hit)
                group = table.filter(mask)
            if batch_format == "pandas":
                out = fn(group.to_pandas())
                pieces.append(pa.Table.from_pandas(out, preserve_index=False))
            elif batch_format == "numpy":
                out = fn({n: group.column(n).to_numpy(zero_copy_only=False) for n in group.column_names})
                pieces.append(pa.table({k: pa.array(v) for k, v in out.items()}))
            else:
                out = fn(group)
                pieces.append(out if isinstance(out, pa.Table) else pa.Table.from_pandas(out, preserve_index=False))
        return from_arrow([pa.concat_tables(pieces, promote_options="default")] if pieces else [table.slice(0, 0)])


# ------------------------------------------------------------------------------------------ Dataset methods
def _install(Dataset) -> None:
    """Attach the extra methods to :class:`raydp_b200.data.Dataset` (kept out of the core module)."""
    import raydp_b200.data as core
    from raydp_b200 import runtime

    def groupby(self, key=None) -> GroupedData:
        return GroupedData(self, key)

    def aggregate(self, *aggs: AggregateFn):
        tables = list(self.iter_tables())
        out = _aggregate_tables(tables, aggs)
        return out

    def _one(self, cls, on, **kw):
        cols = [on] if isinstance(on, str) else (list(on) if on is not None else
                                                 [f.name for f in self.to_arrow().schema
                                                  if pa.types.is_integer(f.type) or pa.types.is_floating(f.type)])
        res = aggregate(self, *[cls(c, **kw) for c in cols])
        values = list(res.values())
        return values[0] if isinstance(on, str) or len(values) == 1 else res

    def sum_(self, on=None, ignore_nulls: bool = True):
        return _one(self, Sum, on)

    def min_(self, on=None, ignore_nulls: bool = True):
        return _one(self, Min, on)

    def max_(self, on=None, ignore_nulls: bool = True):
        return _one(self, Max, on)

    def mean(self, on=None, ignore_nulls: bool = True):
        return _one(self, Mean, on)

    def std(self, on=None, ddof: int = 1, ignore_nulls: bool = True):
        return _one(self, Std, on, ddof=ddof)

    def unique(self, column: str) -> list:
        return Unique(column).combine([Unique(column).of_table(t) for t in self.iter_tables()])

    def flat_map(self, fn: Callable[[Dict[str, Any]], List[Dict[str, Any]]], **_kw):
        def per_block(table: pa.Table, _fn=fn) -> pa.Table:
            rows = [out for r in table.to_pylist() for out in _fn(r)]
            return pa.Table.from_pylist(rows) if rows else table.slice(0, 0)
        return self.map_batches(per_block)

    def rename_columns(self, names: Union[Dict[str, str], List[str]], **_kw):
        def per_block(table: pa.Table, _names=names) -> pa.Table:
            new = [_names.get(c, c) for c in table.column_names] if isinstance(_names, dict) else list(_names)
            return table.rename_columns(new)
        return self.map_batches(per_block)

    def take_batch(self, batch_size: int = 20, batch_format: str = "numpy"):
        for batch in self.iter_batches(batch_size=batch_size, batch_format=batch_format):
            return batch
        raise ValueError("The dataset is empty.")

    def random_sample(self, fraction: float, seed: Optional[int] = None):
        if not 0 <= fraction <= 1:
            raise ValueError("Fraction must be between 0 and 1.")
        base = 0x5eed if seed is None else int(seed)
        kept = []
        for i, table in enumerate(self.iter_tables()):
            rng = np.random.default_rng([base, i])
            kept.append(table.filter(pa.array(rng.random(table.num_rows) < fraction)))
        return core.from_arrow(kept) if kept else self

    def zip_(self, other):
        left, right = self.to_arrow(), other.to_arrow()
        if left.num_rows != right.num_rows:
            raise ValueError("Cannot zip datasets of different number of rows: "
                             f"{left.num_rows}, {right.num_rows}")
        out = left
        for name in right.column_names:
            new = name
            k = 1
            while new in out.column_names:
                new = f"{name}_{k}"
                k += 1
            out = out.append_column(new, right.column(name))
        n = builtins.max(1, self.num_blocks())
        return core.from_arrow(_split(out, n))

    def join(self, ds, join_type: str = "inner", num_partitions: Optional[int] = None, on=None, right_on=None,
             left_suffix: Optional[str] = None, right_suffix: Optional[str] = None, **_kw):
        keys = [on] if isinstance(on, str) else list(on)
        rkeys = keys if right_on is None else ([right_on] if isinstance(right_on, str) else list(right_on))
        how = {"inner": "inner", "left_outer": "left outer", "right_outer": "right outer", "full_outer": "full outer",
               "left_semi": "left semi", "left_anti": "left anti", "right_semi": "right semi",
               "right_anti": "right anti"}[join_type]
        out = self.to_arrow().join(ds.to_arrow(), keys=keys, right_keys=rkeys, join_type=how,
                                   left_suffix=left_suffix, right_suffix=right_suffix, use_threads=False)
        return core.from_arrow(_split(out, builtins.max(1, num_partitions or self.num_blocks())))

    def _write(self, path: str, suffix: str, writer: Callable[[pa.Table, str], None]) -> None:
        os.makedirs(path, exist_ok=True)
        for i, table in enumerate(self.iter_tables()):
            writer(table, os.path.join(path, f"block-{i:05d}.{suffix}"))

    def write_csv(self, path: str, **_kw) -> None:
        import pyarrow.csv as pacsv
        _write(self, path, "csv", lambda t, f: pacsv.write_csv(t, f))

    def write_json(self, path: str, **_kw) -> None:
        import json

        def one(table: pa.Table, f: str) -> None:
            with open(f, "w", encoding="utf-8") as fh:
                for row in table.to_pylist():
                    fh.write(json.dumps(row, default=str) + "\n")
        _write(self, path, "json", one)

    def write_numpy(self, path: str, column: str = "data", **_kw) -> None:
        _write(self, path, "npy", lambda t, f: np.save(f, np.asarray(t.column(column).to_pylist())))

    def to_pandas_refs(self) -> List[Any]:
        return [runtime.put(t.to_pandas()) for t in self.iter_tables()]

    def to_numpy_refs(self, column: Optional[str] = None) -> List[Any]:
        def conv(t: pa.Table):
            if column is not None:
                return t.column(column).to_numpy(zero_copy_only=False)
            return {n: t.column(n).to_numpy(zero_copy_only=False) for n in t.column_names}
        return [runtime.put(conv(t)) for t in self.iter_tables()]

    def input_files(self) -> List[str]:
        return list(getattr(self, "_input_files", []))

    def is_fully_executed(self) -> bool:
        return True

    def fully_executed(self):
        return self.materialize()

    def streaming_split(self, n: int, equal: bool = False, locality_hints=None):
        return [DataIterator(part) for part in self.split(n, equal=equal)]

    def iterator(self) -> "DataIterator":
        return DataIterator(self)

    def iter_tf_batches(self, batch_size: int = 256, **_kw):
        try:
            import tensorflow as tf
        except ImportError as exc:
            raise ImportError("iter_tf_batches needs TensorFlow, which is not installed") from exc
        for batch in self.iter_batches(batch_size=batch_size, batch_format="numpy"):
            yield {k: tf.convert_to_tensor(v) for k, v in batch.items()}

    def to_dask(self, **_kw):
        raise ImportError("Dataset.to_dask needs dask, which is not installed")

    def repeat(self, times: Optional[int] = None):
        if times is None:
            raise NotImplementedError("an endless repeat() pipeline is not supported: pass times=<n> or loop over epochs")
        return self.union(*[self for _ in builtins.range(times - 1)]) if times > 1 else self

    def window(self, *, blocks_per_window: Optional[int] = None, bytes_per_window: Optional[int] = None):
        """Windows of consecutive blocks as separate datasets (the old DatasetPipeline's unit of work)."""
        per = builtins.max(1, blocks_per_window or 1)
        return [core.Dataset(self._blocks[i:i + per], self._schema) for i in builtins.range(0, len(self._blocks), per)]

    def __iter_not_allowed(self):
        raise TypeError("`Dataset` objects aren't iterable. Use iter_rows(), iter_batches() or take_all().")

    methods = dict(groupby=groupby, aggregate=aggregate, sum=sum_, min=min_, max=max_, mean=mean, std=std,
                   unique=unique, flat_map=flat_map, rename_columns=rename_columns, take_batch=take_batch,
                   random_sample=random_sample, zip=zip_, join=join, write_csv=write_csv, write_json=write_json,
                   write_numpy=write_numpy, to_pandas_refs=to_pandas_refs, to_numpy_refs=to_numpy_refs,
                   input_files=input_files, is_fully_executed=is_fully_executed, fully_executed=fully_executed,
                   streaming_split=streaming_split, iterator=iterator, iter_tf_batches=iter_tf_batches,
                   to_dask=to_dask, repeat=repeat, window=window, __iter__=__iter_not_allowed)
    for name, fn in methods.items():
        if name not in Dataset.__dict__:
            setattr(Dataset, name, fn)


def _split(table: pa.Table, n: int) -> List[pa.Table]:
    n = builtins.max(1, builtins.min(n, builtins.max(1, table.num_rows)))
    per, rem = divmod(table.num_rows, n)
    out, off = [], 0
    for i in builtins.range(n):
        size = per + (1 if i < rem else 0)
        out.append(table.slice(off, size))
        off += size
    return out


class DataIterator:
    """``ray.data.DataIterator``: what ``get_dataset_shard`` / ``streaming_split`` hand to a training
    worker -- iteration only, no transformations."""

    def __init__(self, ds):
        self._ds = ds

    def iter_batches(self, **kw):
        return self._ds.iter_batches(**kw)

    def iter_torch_batches(self, **kw):
        return self._ds.iter_torch_batches(**kw)

    def iter_rows(self):
        return self._ds.iter_rows()

    def to_torch(self, **kw):
        return self._ds.to_torch(**kw)

    def materialize(self):
        return self._ds.materialize()

    def stats(self) -> str:
        return self._ds.stats()

    def schema(self):
        return self._ds.schema()

    def count(self) -> int:
        return self._ds.count()


class DataContext:
    """``ray.data.DataContext.get_current()``: a bag of settings programs set and read back."""

    _current: Optional["DataContext"] = None

    def __init__(self):
        self.target_max_block_size = 128 * 1024 * 1024
        self.target_min_block_size = 1 * 1024 * 1024
        self.enable_progress_bars = False
        self.use_push_based_shuffle = False
        self.execution_options = type("ExecutionOptions", (), {"preserve_order": True, "locality_with_output": False,
                                                               "verbose_progress": False})()

    @classmethod
    def get_current(cls) -> "DataContext":
        if cls._current is None:
            cls._current = cls()
        return cls._current
