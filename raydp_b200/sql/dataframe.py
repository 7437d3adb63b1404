"""PySpark-shaped DataFrame over partitioned Arrow tables.

Parity: the DataFrame surface the reference exercises: ``withColumn,
filter, select, drop, randomSplit, repartition, sortWithinPartitions, cache/persist/unpersist,
count, take, collect, show, schema, rdd.getNumPartitions, groupBy().count(), fillna, join,
withColumnRenamed, agg, mapInPandas, write.parquet`` (reference ``examples/data_process.py``,
``python/raydp/utils.py:77``, ``spark/dataset.py:177-180,268``, ``torch/estimator.py:371``).
Narrow operators extend the plan; wide ones run a two-stage job through the session backend.
"""
from __future__ import annotations

import uuid
from typing import Any, Dict, Iterable, List, Optional, Sequence, Tuple, Union

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

from raydp_b200.sql import types as T
from raydp_b200.sql.column import Column, as_column
from raydp_b200.sql.plan import LocalSource, Plan, RefSource


class StorageLevel:
    """Accepted for API compatibility; blocks always live in executor memory / shared memory."""

    def __init__(self, name: str):
        self.name = name

    def __repr__(self):
        return f"StorageLevel.{self.name}"


for _n in ("MEMORY_ONLY", "MEMORY_AND_DISK", "DISK_ONLY", "MEMORY_ONLY_2", "MEMORY_AND_DISK_2",
           "OFF_HEAP", "MEMORY_AND_DISK_DESER"):
    setattr(StorageLevel, _n, StorageLevel(_n))


class DataFrame:
    def __init__(self, session, plan: Plan):
        self._session = session
        self._plan = plan
        self._cache_pending = False
        self._cached_refs: Optional[list] = None
        self._lineage: Optional[Plan] = None
        self._schema_cache: Optional[pa.Schema] = None

    # ------------------------------------------------------------------ plumbing
    @property
    def sparkSession(self):
        return self._session

    @property
    def sql_ctx(self):
        return _SQLContextShim(self._session)

    def _derive(self, op: tuple) -> "DataFrame":
        self._materialise_cache()
        return DataFrame(self._session, self._plan.with_op(op))

    def _materialise_cache(self) -> None:
        if not self._cache_pending:
            return
        self._cache_pending = False
        results = self._session._backend.run(self._plan, "put", {})
        schema = self._plan.schema()
        self._lineage = self._plan
        self._cached_refs = [[r[0]] for r in results]
        self._plan = Plan(RefSource(self._cached_refs, schema))

    def _run(self, action: str, args: Optional[dict] = None, plan: Optional[Plan] = None) -> list:
        self._materialise_cache()
        return self._session._backend.run(plan or self._plan, action, args or {})

    def _arrow_schema(self) -> pa.Schema:
        if self._schema_cache is None:
            self._schema_cache = self._plan.schema()
        return self._schema_cache

    # ------------------------------------------------------------------ metadata
    @property
    def schema(self) -> T.StructType:
        return T.from_arrow_schema(self._arrow_schema())

    @property
    def columns(self) -> List[str]:
        return list(self._arrow_schema().names)

    @property
    def dtypes(self) -> List[Tuple[str, str]]:
        return [(f.name, f.dataType.simpleString()) for f in self.schema]

    def printSchema(self) -> None:
        print("root")
        for f in self.schema:
            print(f" |-- {f.name}: {f.dataType.simpleString()} (nullable = {str(f.nullable).lower()})")

    @property
    def rdd(self):
        from raydp_b200.sql.rdd import DataFrameRDD
        return DataFrameRDD(self)

    def __getitem__(self, item):
        if isinstance(item, str):
            if item not in self.columns:
                raise KeyError(f"cannot resolve column '{item}' given input columns {self.columns}")
            return Column("col", (item,))
        if isinstance(item, Column):
            return self.filter(item)
        if isinstance(item, (list, tuple)):
            return self.select(*item)
        raise TypeError(f"unexpected item type: {type(item)}")

    def __getattr__(self, name: str):
        if name.startswith("_"):
            raise AttributeError(name)
        if name in self.columns:
            return Column("col", (name,))
        raise AttributeError(f"'DataFrame' object has no attribute '{name}'")

    def __repr__(self):
        return "DataFrame[" + ", ".join(f"{n}: {t}" for n, t in self.dtypes) + "]"

    # ------------------------------------------------------------------ narrow transforms
    def withColumn(self, name: str, col: Column) -> "DataFrame":
        if not isinstance(col, Column):
            raise TypeError("col should be Column")
        if col._kind in ("window",):
            return self._with_window_column(name, col)
        if col._kind in ("posexplode", "explode"):
            raise NotImplementedError("use select(..., posexplode(...)) for generators")
        return self._derive(("with_column", name, col))

    def withColumns(self, cols: Dict[str, Column]) -> "DataFrame":
        df = self
        for k, v in cols.items():
            df = df.withColumn(k, v)
        return df

    def withColumnRenamed(self, existing: str, new: str) -> "DataFrame":
        return self._derive(("rename", existing, new))

    def toDF(self, *names: str) -> "DataFrame":
        if len(names) != len(self.columns):
            raise ValueError("number of names does not match number of columns")
        return self._derive(("rename_all", tuple(names)))

    def filter(self, condition: Union[Column, str]) -> "DataFrame":
        if isinstance(condition, str):
            from raydp_b200.sql.sqlparser import parse_expression
            condition = parse_expression(condition)
        return self._derive(("filter", condition))

    where = filter

    def select(self, *cols) -> "DataFrame":
        if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
            cols = tuple(cols[0])
        exprs = [as_column(c) for c in cols]
        gens = [e for e in exprs if e._kind in ("posexplode", "explode")]
        if gens:
            keep = [e._args[0] for e in exprs if e._kind == "col"]
            df = self._derive(("posexplode", gens[0]._args[0], keep))
            if gens[0]._kind == "explode":
                df = df.drop("pos")
            return df
        if any(e._kind == "agg" for e in exprs):
            return self.agg(*exprs)
        return self._derive(("select", exprs))

    def selectExpr(self, *exprs) -> "DataFrame":
        from raydp_b200.sql.sqlparser import parse_expression
        if len(exprs) == 1 and isinstance(exprs[0], (list, tuple)):
            exprs = tuple(exprs[0])
        cols: list = []
        for e in exprs:
            cols.extend(self.columns if e.strip() == "*" else [parse_expression(e)])
        return self.select(*cols)

    def createOrReplaceTempView(self, name: str) -> None:
        self._session.catalog.register_view(name, self)

    createTempView = createOrReplaceTempView
    registerTempTable = createOrReplaceTempView

    def drop(self, *cols) -> "DataFrame":
        names = [c if isinstance(c, str) else c._output_name() for c in cols]
        return self._derive(("drop", names))

    def fillna(self, value, subset: Optional[Union[str, Sequence[str]]] = None) -> "DataFrame":
        if isinstance(subset, str):
            subset = [subset]
        return self._derive(("fillna", value, list(subset) if subset is not None else None))

    @property
    def na(self):
        return _NAFunctions(self)

    def dropna(self, how: str = "any", thresh=None, subset=None) -> "DataFrame":
        if isinstance(subset, str):
            subset = [subset]
        return self._derive(("dropna", list(subset) if subset else None))

    def sortWithinPartitions(self, *cols, ascending: Union[bool, List[bool]] = True) -> "DataFrame":
        return self._derive(("sort_within", _sort_keys(cols, ascending)))

    def sample(self, withReplacement=None, fraction=None, seed=None) -> "DataFrame":
        if isinstance(withReplacement, float) and fraction is None:
            fraction, withReplacement = withReplacement, False
        seed = int(seed) if seed is not None else int(np.random.SeedSequence().entropy % (2 ** 31))
        return self._derive(("sample", float(fraction), seed))

    def randomSplit(self, weights: List[float], seed: Optional[int] = None) -> List["DataFrame"]:
        if any(w < 0 for w in weights):
            raise ValueError("Weights must be positive. Found weight value: %s" % min(weights))
        seed = int(seed) if seed is not None else int(np.random.SeedSequence().entropy % (2 ** 31))
        total = float(sum(weights))
        bounds = np.concatenate([[0.0], np.cumsum([w / total for w in weights])])
        bounds[-1] = 1.0 + 1e-12
        return [self._derive(("split", float(bounds[i]), float(bounds[i + 1]), seed))
                for i in range(len(weights))]

    def mapInPandas(self, func, schema) -> "DataFrame":
        if isinstance(schema, str):
            schema = T.parse_ddl(schema)
        arrow_schema = schema.to_arrow_schema() if isinstance(schema, T.StructType) else schema
        return self._derive(("map_pandas", func, arrow_schema))

    def mapInArrow(self, func, schema=None) -> "DataFrame":
        return self._derive(("map_arrow", func))

    def distinct(self) -> "DataFrame":
        keys = self.columns
        return self.repartition(self.rdd.getNumPartitions(), *keys)._derive(("distinct",))

    def dropDuplicates(self, subset=None) -> "DataFrame":
        if subset:
            keys = [subset] if isinstance(subset, str) else list(subset)
            return self.repartition(self.rdd.getNumPartitions(), *keys)._derive(("distinct_subset", tuple(keys)))
        return self.distinct()

    drop_duplicates = dropDuplicates

    def unionByName(self, other: "DataFrame", allowMissingColumns: bool = False) -> "DataFrame":
        mine, theirs = self.columns, other.columns
        if allowMissingColumns:
            from raydp_b200.sql import functions as F
            for c in mine:
                if c not in theirs:
                    other = other.withColumn(c, F.lit(None))
            for c in theirs:
                if c not in mine:
                    self = self.withColumn(c, F.lit(None))     # noqa: PLW0642
            mine = self.columns
        elif set(mine) != set(theirs):
            raise ValueError(f"unionByName: column sets differ: {mine} vs {theirs}")
        return self.union(other.select(*mine))

    def limit(self, n: int) -> "DataFrame":
        table = self._collect_table(limit=n)
        return self._session._from_tables([table.slice(0, n)])

    # ------------------------------------------------------------------ wide transforms
    def repartition(self, numPartitions, *cols) -> "DataFrame":
        if not isinstance(numPartitions, int):
            cols = (numPartitions,) + cols
            numPartitions = self.rdd.getNumPartitions()
        keys = [c if isinstance(c, str) else c._output_name() for c in cols]
        schema = self._arrow_schema()
        import uuid
        pieces = self._run("shuffle_map", {"n_out": numPartitions, "keys": keys or None,
                                           "shuffle_id": uuid.uuid4().hex[:12]})
        refs = [[row[d] for row in pieces if row[d] is not None] for d in range(numPartitions)]
        out = DataFrame(self._session, Plan(RefSource(refs, schema)))
        out._cached_refs = refs
        self._session._track_refs([r for part in refs for r in part])
        return out

    def coalesce(self, numPartitions: int) -> "DataFrame":
        if numPartitions >= self.rdd.getNumPartitions():
            return self
        return self.repartition(numPartitions)

    def groupBy(self, *cols) -> "GroupedData":
        if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
            cols = tuple(cols[0])
        return GroupedData(self, [c if isinstance(c, str) else c._output_name() for c in cols])

    groupby = groupBy

    def agg(self, *exprs) -> "DataFrame":
        return GroupedData(self, []).agg(*exprs)

    def join(self, other: "DataFrame", on=None, how: str = "inner") -> "DataFrame":
        """Broadcast-hash join: the right side is collected once and shipped with the plan.  ``on`` is a
        column name, a list of names, or an equality expression ``left.a == right.b`` (conjunctions
        with ``&`` allowed).  Right / full outer joins first gather the left side into one partition
        (unmatched right rows must be emitted exactly once)."""
        how = how.lower()
        if on is None or how == "cross":
            return self.crossJoin(other)
        if isinstance(on, Column):
            pairs = _equi_join_pairs(on)
            keys, right_keys = [p[0] for p in pairs], [p[1] for p in pairs]
        else:
            keys = [on] if isinstance(on, str) else [k if isinstance(k, str) else k._output_name() for k in on]
            right_keys = keys
        left = self
        if how in ("right", "right_outer", "rightouter", "outer", "full", "full_outer", "fullouter"):
            left = self.coalesce(1)
        right = other._collect_table()
        how = {"rightouter": "right_outer", "fullouter": "full_outer"}.get(how, how)
        return left._derive(("join", right, keys, how, right_keys))

    def crossJoin(self, other: "DataFrame") -> "DataFrame":
        from raydp_b200.sql import functions as F
        key = "__cross_key"
        joined = self.withColumn(key, F.lit(1))._derive(
            ("join", other.withColumn(key, F.lit(1))._collect_table(), [key], "inner", [key]))
        return joined.drop(key)

    def union(self, other: "DataFrame") -> "DataFrame":
        tables = self._run("collect") + other._run("collect")
        schema = self._arrow_schema()
        tables = [t.rename_columns(schema.names).cast(schema, safe=False) for t in tables]
        return self._session._from_tables(tables)

    unionAll = union

    def orderBy(self, *cols, ascending: Union[bool, List[bool]] = True) -> "DataFrame":
        table = self._collect_table().sort_by(_sort_keys(cols, ascending))
        n = max(1, self.rdd.getNumPartitions())
        return self._session._from_tables(_split_table(table, n))

    sort = orderBy

    def _with_window_column(self, name: str, col: Column) -> "DataFrame":
        fn, spec = col._args
        if fn._kind != "rank_fn" or fn._args[0] != "row_number":
            raise NotImplementedError("only row_number() window functions are supported")
        part_keys = [c._output_name() for c in spec._partition_by]
        table = self._collect_table()
        order = [(c._args[0]._output_name() if c._kind == "sort" else c._output_name(),
                  c._args[1] if c._kind == "sort" else "ascending") for c in spec._order_by]
        idx = pc.sort_indices(table, sort_keys=[(k, "ascending") for k in part_keys] + order)
        sorted_t = table.take(idx)
        n = sorted_t.num_rows
        if part_keys and n:
            change = np.zeros(n, dtype=bool)
            change[0] = True
            for k in part_keys:
                v = sorted_t.column(k).combine_chunks()
                neq = pc.not_equal(v.slice(1), v.slice(0, n - 1)).to_numpy(zero_copy_only=False)
                change[1:] |= neq.astype(bool)
            starts = np.maximum.accumulate(np.where(change, np.arange(n), 0))
            rn = np.arange(n) - starts + 1
        else:
            rn = np.arange(1, n + 1)
        sorted_t = sorted_t.append_column(name, pa.array(rn.astype(np.int32)))
        return self._session._from_tables(_split_table(sorted_t, max(1, self.rdd.getNumPartitions())))

    # ------------------------------------------------------------------ persistence
    def cache(self) -> "DataFrame":
        return self.persist()

    def persist(self, storageLevel: Optional[StorageLevel] = None) -> "DataFrame":
        if self._cached_refs is None:
            self._cache_pending = True
        return self

    def unpersist(self, blocking: bool = False) -> "DataFrame":
        self._cache_pending = False
        if self._cached_refs is not None and self._lineage is not None:
            from raydp_b200 import runtime
            runtime.free([r for part in self._cached_refs for r in part])
            self._plan = self._lineage
            self._cached_refs = None
            self._lineage = None
        return self

    @property
    def is_cached(self) -> bool:
        return self._cache_pending or self._cached_refs is not None

    @property
    def storageLevel(self) -> StorageLevel:
        return StorageLevel.MEMORY_AND_DISK if self.is_cached else StorageLevel("NONE")

    # ------------------------------------------------------------------ actions
    def count(self) -> int:
        return int(sum(self._run("count")))

    def _collect_table(self, limit: Optional[int] = None) -> pa.Table:
        schema = self._arrow_schema()
        if limit is not None:
            tables = self._run("take", {"n": limit})
        else:
            tables = self._run("collect")
        tables = [t for t in tables if t.num_rows] or [schema.empty_table()]
        return pa.concat_tables([t.cast(schema, safe=False) if t.schema != schema else t
                                 for t in tables]) if len(tables) > 1 else tables[0]

    def toArrow(self) -> pa.Table:
        return self._collect_table()

    def toPandas(self):
        return self._collect_table().to_pandas()

    def collect(self) -> List[T.Row]:
        table = self._collect_table()
        names = table.column_names
        cols = [c.to_pylist() for c in table.columns]
        return [T.Row._make(names, vals) for vals in zip(*cols)] if cols else []

    def take(self, num: int) -> List[T.Row]:
        table = self._collect_table(limit=num).slice(0, num)
        names = table.column_names
        cols = [c.to_pylist() for c in table.columns]
        return [T.Row._make(names, vals) for vals in zip(*cols)] if cols else []

    def head(self, n: Optional[int] = None):
        rows = self.take(1 if n is None else n)
        return (rows[0] if rows else None) if n is None else rows

    def first(self):
        return self.head()

    def isEmpty(self) -> bool:
        return len(self.take(1)) == 0

    def show(self, n: int = 20, truncate: bool = True, vertical: bool = False) -> None:
        table = self._collect_table(limit=n + 1)
        more = table.num_rows > n
        table = table.slice(0, n)
        names = table.column_names
        cells = [[_fmt(v, truncate) for v in c.to_pylist()] for c in table.columns]
        widths = [max([len(nm)] + [len(v) for v in col]) for nm, col in zip(names, cells)]
        line = "+" + "+".join("-" * w for w in widths) + "+"
        print(line)
        print("|" + "|".join(nm.rjust(w) for nm, w in zip(names, widths)) + "|")
        print(line)
        for i in range(table.num_rows):
            print("|" + "|".join(col[i].rjust(w) for col, w in zip(cells, widths)) + "|")
        print(line)
        if more:
            print(f"only showing top {n} row{'s' if n != 1 else ''}")

    def foreachPartition(self, f) -> None:
        def run(table, _f=f):
            names = table.column_names
            cols = [c.to_pylist() for c in table.columns]
            _f(iter([T.Row._make(names, vals) for vals in zip(*cols)]))
            return None
        self._run("py_partition", {"fn": run})

    def describe(self, *cols) -> "DataFrame":
        table = self._collect_table()
        names = list(cols) or [f.name for f in table.schema
                               if pa.types.is_integer(f.type) or pa.types.is_floating(f.type)]
        rows = {"summary": ["count", "mean", "stddev", "min", "max"]}
        for nme in names:
            c = table.column(nme)
            rows[nme] = [str(pc.count(c).as_py()), str(pc.mean(c).as_py()),
                         str(pc.stddev(c, ddof=1).as_py()), str(pc.min(c).as_py()),
                         str(pc.max(c).as_py())]
        return self._session._from_tables([pa.table(rows)])

    @property
    def write(self) -> "DataFrameWriter":
        return DataFrameWriter(self)

    def to_pandas_on_spark(self, index_col=None):
        from raydp_b200.sql import pandas_api
        return pandas_api.DataFrame(self)

    pandas_api = to_pandas_on_spark
    to_koalas = to_pandas_on_spark


class _SQLContextShim:
    def __init__(self, session):
        self.sparkSession = session


class _NAFunctions:
    def __init__(self, df: DataFrame):
        self._df = df

    def fill(self, value, subset=None):
        return self._df.fillna(value, subset)

    def drop(self, how="any", thresh=None, subset=None):
        return self._df.dropna(how, thresh, subset)


def _fmt(v, truncate: bool) -> str:
    s = "null" if v is None else str(v)
    return s[:17] + "..." if truncate and len(s) > 20 else s


def _equi_join_pairs(cond: Column) -> List[Tuple[str, str]]:
    """``(a == b) & (c == d)`` -> ``[(a, b), (c, d)]``; anything else is not an equi-join."""
    if cond._kind == "binop" and cond._args[0] == "&":
        return _equi_join_pairs(cond._args[1]) + _equi_join_pairs(cond._args[2])
    if cond._kind == "binop" and cond._args[0] == "==" and all(
            isinstance(a, Column) and a._kind == "col" for a in cond._args[1:3]):
        return [(cond._args[1]._args[0], cond._args[2]._args[0])]
    raise NotImplementedError("join conditions must be equalities between columns (optionally combined with &)")


def _sort_keys(cols, ascending) -> List[Tuple[str, str]]:
    if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
        cols = tuple(cols[0])
    asc = ascending if isinstance(ascending, (list, tuple)) else [ascending] * len(cols)
    keys = []
    for c, a in zip(cols, asc):
        if isinstance(c, Column) and c._kind == "sort":
            keys.append((c._args[0]._output_name(), c._args[1]))
        else:
            name = c if isinstance(c, str) else c._output_name()
            keys.append((name, "ascending" if a else "descending"))
    return keys


def _split_table(table: pa.Table, n: int) -> List[pa.Table]:
    n = max(1, min(n, max(1, table.num_rows)))
    per, rem = divmod(table.num_rows, n)
    out, off = [], 0
    for i in range(n):
        size = per + (1 if i < rem else 0)
        out.append(table.slice(off, size))
        off += size
    return out


class GroupedData:
    def __init__(self, df: DataFrame, keys: List[str]):
        self._df = df
        self._keys = keys

    def count(self) -> DataFrame:
        return self.agg(Column("agg", ("count", None), "count"))

    def _simple(self, fn: str, cols) -> DataFrame:
        if not cols:
            schema = self._df._arrow_schema()
            cols = [f.name for f in schema if f.name not in self._keys and
                    (pa.types.is_integer(f.type) or pa.types.is_floating(f.type))]
        spark_fn = {"mean": "avg"}.get(fn, fn)
        return self.agg(*[Column("agg", (fn, as_column(c)), f"{spark_fn}({c})") for c in cols])

    def sum(self, *cols): return self._simple("sum", cols)
    def avg(self, *cols): return self._simple("mean", cols)
    mean = avg
    def min(self, *cols): return self._simple("min", cols)
    def max(self, *cols): return self._simple("max", cols)

    def agg(self, *exprs) -> DataFrame:
        if len(exprs) == 1 and isinstance(exprs[0], dict):
            exprs = tuple(Column("agg", ({"avg": "mean"}.get(fn, fn), None if c == "*" else as_column(c)),
                                 f"{fn}({c})") for c, fn in exprs[0].items())
        df = self._df
        aggs: List[Tuple[str, Optional[str]]] = []
        names: List[str] = []
        for i, e in enumerate(exprs):
            if e._kind != "agg":
                raise TypeError("agg() expects aggregate expressions such as count('*')")
            fn, target = e._args
            if target is None:
                col_name = None
            elif target._kind == "col":
                col_name = target._args[0]
            else:
                col_name = f"__agg_in_{i}"
                df = df.withColumn(col_name, target)
            aggs.append((fn, col_name))
            default = f"{ {'mean': 'avg'}.get(fn, fn) }({'1' if target is None else target._output_name()})"
            names.append(e._name or default.replace(" ", ""))
        partials = df._run("agg_partial", {"keys": self._keys, "aggs": aggs})
        merged = pa.concat_tables(partials, promote_options="default")
        result = _final_agg(merged, self._keys, aggs, names)
        return df._session._from_tables([result])


def _final_agg(merged: pa.Table, keys: List[str], aggs, names) -> pa.Table:
    """Second aggregation stage: merge per-partition partial states."""
    def grouped(specs):
        if keys:
            return merged.group_by(keys, use_threads=False).aggregate(specs)
        return pa.TableGroupBy(merged, [], use_threads=False).aggregate(specs)

    merge_specs = []
    for fn, col in aggs:
        if fn == "count":
            merge_specs.append(("count_all" if col is None else f"{col}_count", "sum"))
        elif fn == "sum":
            merge_specs.append((f"{col}_sum", "sum"))
        elif fn in ("min", "max"):
            merge_specs.append((f"{col}_{fn}", fn))
        elif fn == "mean":
            merge_specs.extend([(f"{col}_sum", "sum"), (f"{col}_count", "sum")])
        elif fn in ("stddev", "variance"):
            merge_specs.extend([(f"{col}_sum", "sum"), (f"__sq_{col}_sum", "sum"), (f"{col}_count", "sum")])
        elif fn == "first":
            merge_specs.append((f"{col}_first", "first"))
        elif fn == "count_distinct":
            merge_specs.append((f"{col}_distinct", "list"))
    uniq = list(dict.fromkeys(merge_specs))
    g = grouped(uniq)
    out_cols, out_names = [g.column(k) for k in keys], list(keys)
    for (fn, col), name in zip(aggs, names):
        if fn == "count":
            src = "count_all_sum" if col is None else f"{col}_count_sum"
            out_cols.append(pc.cast(g.column(src), pa.int64()))
        elif fn == "sum":
            out_cols.append(g.column(f"{col}_sum_sum"))
        elif fn in ("min", "max"):
            out_cols.append(g.column(f"{col}_{fn}_{fn}"))
        elif fn == "mean":
            out_cols.append(pc.divide(pc.cast(g.column(f"{col}_sum_sum"), pa.float64()),
                                      pc.cast(g.column(f"{col}_count_sum"), pa.float64())))
        elif fn in ("stddev", "variance"):
            s = pc.cast(g.column(f"{col}_sum_sum"), pa.float64())
            sq = pc.cast(g.column(f"__sq_{col}_sum_sum"), pa.float64())
            cnt = pc.cast(g.column(f"{col}_count_sum"), pa.float64())
            var = pc.divide(pc.subtract(sq, pc.divide(pc.multiply(s, s), cnt)), pc.subtract(cnt, 1.0))
            out_cols.append(pc.sqrt(var) if fn == "stddev" else var)
        elif fn == "first":
            out_cols.append(g.column(f"{col}_first_first"))
        elif fn == "count_distinct":
            lists = g.column(f"{col}_distinct_list").to_pylist()
            out_cols.append(pa.array([len({v for sub in row for v in sub if v is not None})
                                      for row in lists], type=pa.int64()))
        out_names.append(name)
    return pa.table(out_cols, names=out_names)


class DataFrameWriter:
    def __init__(self, df: DataFrame):
        self._df = df
        self._mode = "errorifexists"
        self._options: Dict[str, Any] = {}

    def mode(self, saveMode: str) -> "DataFrameWriter":
        self._mode = (saveMode or "errorifexists").lower()
        return self

    def option(self, key: str, value) -> "DataFrameWriter":
        self._options[key.lower()] = value
        return self

    def format(self, fmt: str) -> "DataFrameWriter":
        self._options["__format"] = fmt
        return self

    def _prepare(self, path: str) -> Optional[str]:
        import os
        import shutil
        from raydp_b200.sql.plan import strip_scheme
        path = strip_scheme(path)
        if os.path.exists(path) and (os.path.isfile(path) or os.listdir(path)):
            if self._mode in ("overwrite",):
                shutil.rmtree(path, ignore_errors=True)
            elif self._mode in ("ignore",):
                return None
            elif self._mode in ("append",):
                pass
            else:
                raise FileExistsError(f"path {path} already exists.")
        os.makedirs(path, exist_ok=True)
        return path

    def parquet(self, path: str, mode: Optional[str] = None, compression: Optional[str] = None) -> None:
        import os
        if mode:
            self.mode(mode)
        target = self._prepare(path)
        if target is None:
            return
        compression = compression or self._options.get("compression")
        self._df._run("write_parquet", {"path": target, "compression": compression,
                                        "job": uuid.uuid4().hex[:8]})
        open(os.path.join(target, "_SUCCESS"), "w").close()

    def saveAsTable(self, name: str, format: Optional[str] = None, mode: Optional[str] = None, **_kw) -> None:  # noqa: A002
        """Persist as a parquet table under the warehouse directory (see ``sqlparser.Catalog``)."""
        self._df._session.catalog.save_table(self._df, name, (mode or self._mode))

    def insertInto(self, name: str, overwrite: bool = False) -> None:
        self._df._session.catalog.save_table(self._df, name, "overwrite" if overwrite else "append")

    def csv(self, path: str, mode: Optional[str] = None, header=None, **_kw) -> None:
        import os
        if mode:
            self.mode(mode)
        target = self._prepare(path)
        if target is None:
            return
        hdr = header if header is not None else self._options.get("header", False)
        hdr = str(hdr).lower() == "true" if not isinstance(hdr, bool) else hdr
        self._df._run("write_csv", {"path": target, "header": hdr, "job": uuid.uuid4().hex[:8]})
        open(os.path.join(target, "_SUCCESS"), "w").close()

    def save(self, path: str) -> None:
        fmt = self._options.get("__format", "parquet")
        getattr(self, fmt)(path)
