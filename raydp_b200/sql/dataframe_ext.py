"""The rest of the ``pyspark.sql.DataFrame`` surface, as a mixin of :class:`raydp_b200.sql.dataframe.DataFrame`.

The reference hands users a real Spark DataFrame (``raydp.init_spark`` returns a SparkSession,
reference ``python/raydp/context.py:140-206``), so code moving over calls whatever pyspark offers:
set operations, ``stat`` helpers, ``summary``, ``melt``, ``rollup`` / ``cube``, ``replace``, ``toJSON`` ...
These are built from the engine's primitives (select / filter / join / group-by over Arrow tables).
"""
from __future__ import annotations

import itertools
import json
from typing import Any, Callable, Dict, Iterator, List, Optional, Sequence, Tuple, Union

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

from raydp_b200.sql import types as T
from raydp_b200.sql.column import Column, as_column


def _row_keys(table: pa.Table) -> List[tuple]:
    from raydp_b200.sql.dataframe import _hashable
    cols = [c.to_pylist() for c in table.columns]
    return [tuple(_hashable(v) for v in row) for row in zip(*cols)] if cols else [()] * table.num_rows


class DataFrameExtras:
    # ------------------------------------------------------------------ identity / plan
    def alias(self, alias: str):
        out = self.select(*self.columns)
        out._alias = alias
        return out

    def hint(self, name: str, *parameters):
        return self

    def checkpoint(self, eager: bool = True):
        return self.persist()

    localCheckpoint = checkpoint

    def isLocal(self) -> bool:
        return False

    @property
    def isStreaming(self) -> bool:
        return False

    def explain(self, extended=None, mode: Optional[str] = None) -> None:
        plan = self._plan
        print("== Physical Plan ==")
        for depth, op in enumerate(reversed(plan.ops)):
            detail = ", ".join(repr(a)[:60] for a in op[1:3])
            print(f"{'   ' * depth}{'+- ' if depth else ''}{op[0]}({detail})")
        print(f"{'   ' * len(plan.ops)}+- {type(plan.source).__name__}[{plan.source.num_partitions} partitions]")

    def sameSemantics(self, other) -> bool:
        return self.semanticHash() == other.semanticHash()

    def semanticHash(self) -> int:
        return hash((id(self._plan.source), repr(self._plan.ops)))

    def inputFiles(self) -> List[str]:
        return list(getattr(self._plan.source, "files", []) or [])

    def withMetadata(self, columnName: str, metadata: dict):
        if columnName not in self.columns:
            raise KeyError(columnName)
        return self

    def observe(self, observation, *exprs):
        return self

    def transform(self, func: Callable, *args, **kwargs):
        out = func(self, *args, **kwargs)
        if not hasattr(out, "_plan"):
            raise TypeError(f"transform: the function returned {type(out).__name__}, not a DataFrame")
        return out

    def colRegex(self, colName: str) -> List[Column]:
        import re
        rx = re.compile(colName.strip("`"))
        return [Column("col", (c,)) for c in self.columns if rx.fullmatch(c)]

    def withColumnsRenamed(self, colsMap: Dict[str, str]):
        df = self
        for old, new in colsMap.items():
            df = df.withColumnRenamed(old, new)
        return df

    def createGlobalTempView(self, name: str) -> None:
        self._session.catalog.register_view(f"global_temp.{name}", self)
        self._session.catalog.register_view(name, self)

    createOrReplaceGlobalTempView = createGlobalTempView

    # ------------------------------------------------------------------ rows out
    def tail(self, num: int) -> List[T.Row]:
        rows = self.collect()
        return rows[-num:] if num > 0 else []

    def toLocalIterator(self, prefetchPartitions: bool = False) -> Iterator[T.Row]:
        return iter(self.collect())

    def foreach(self, f: Callable) -> None:
        self.foreachPartition(lambda rows: [f(r) for r in rows])

    def toJSON(self, use_unicode: bool = True):
        from raydp_b200.sql.fnlib import _jsonable
        names = self.columns
        table = self._collect_table()
        cols = [c.to_pylist() for c in table.columns]
        lines = [json.dumps({k: _jsonable(v) for k, v in zip(names, row) if v is not None}, separators=(",", ":"))
                 for row in zip(*cols)]
        from raydp_b200.sql.rdd import RDD
        n = max(1, self.rdd.getNumPartitions())
        return RDD(self._session, _chunks(lines, n))

    def offset(self, num: int):
        table = self._collect_table()
        return self._session._from_tables([table.slice(num)])

    # ------------------------------------------------------------------ set operations
    def _set_op(self, other, keep: Callable[[int, int], int]):
        """Multiset arithmetic on whole rows: ``keep(count_left, count_right)`` copies of each distinct row."""
        left, right = self._collect_table(), other._collect_table()
        lkeys, rkeys = _row_keys(left), _row_keys(right)
        rcount: Dict[tuple, int] = {}
        for k in rkeys:
            rcount[k] = rcount.get(k, 0) + 1
        lcount: Dict[tuple, int] = {}
        for k in lkeys:
            lcount[k] = lcount.get(k, 0) + 1
        emitted: Dict[tuple, int] = {}
        take = []
        for i, k in enumerate(lkeys):
            quota = keep(lcount[k], rcount.get(k, 0))
            if emitted.get(k, 0) < quota:
                emitted[k] = emitted.get(k, 0) + 1
                take.append(i)
        return self._session._from_tables([left.take(pa.array(take, type=pa.int64()))])

    def subtract(self, other):
        return self._set_op(other, lambda l, r: 1 if r == 0 else 0)

    def exceptAll(self, other):
        return self._set_op(other, lambda l, r: max(l - r, 0))

    def intersect(self, other):
        return self._set_op(other, lambda l, r: 1 if r > 0 else 0)

    def intersectAll(self, other):
        return self._set_op(other, lambda l, r: min(l, r))

    def unionAll(self, other):
        return self.union(other)

    # ------------------------------------------------------------------ reshaping
    def melt(self, ids, values, variableColumnName: str, valueColumnName: str):
        ids = [ids] if isinstance(ids, (str, Column)) else list(ids or [])
        id_names = [c if isinstance(c, str) else c._output_name() for c in ids]
        if values is None:
            values = [c for c in self.columns if c not in id_names]
        values = [values] if isinstance(values, (str, Column)) else list(values)
        value_names = [c if isinstance(c, str) else c._output_name() for c in values]
        if not value_names:
            raise ValueError("[UNPIVOT_REQUIRES_VALUE_COLUMNS] at least one value column is required")
        table = self._collect_table()
        n, k = table.num_rows, len(value_names)
        types = [table.schema.field(v).type for v in value_names]
        common = types[0]
        for t in types[1:]:
            if t != common:
                if (pa.types.is_integer(common) or pa.types.is_floating(common)) and \
                        (pa.types.is_integer(t) or pa.types.is_floating(t)):
                    common = pa.float64() if pa.types.is_floating(common) or pa.types.is_floating(t) else pa.int64()
                else:
                    raise TypeError("[UNPIVOT_VALUE_DATA_TYPE_MISMATCH] value columns must share a type")
        order = (np.arange(n, dtype=np.int64)[:, None] + np.arange(k, dtype=np.int64)[None, :] * n).reshape(-1)
        stacked = pa.concat_arrays([table.column(v).combine_chunks().cast(common) for v in value_names]) \
            if n else pa.array([], type=common)
        cols = [table.column(c).combine_chunks().take(pa.array(np.repeat(np.arange(n, dtype=np.int64), k)))
                for c in id_names]
        cols.append(pa.array(np.tile(np.array(value_names, dtype=object), n), type=pa.string()))
        cols.append(stacked.take(pa.array(order)))
        return self._session._from_tables([pa.table(cols, names=id_names + [variableColumnName, valueColumnName])])

    unpivot = melt

    def replace(self, to_replace, value=None, subset=None):
        if isinstance(to_replace, dict):
            mapping = dict(to_replace)
        elif isinstance(to_replace, (list, tuple)):
            vals = value if isinstance(value, (list, tuple)) else [value] * len(to_replace)
            mapping = dict(zip(to_replace, vals))
        else:
            mapping = {to_replace: value}
        subset = [subset] if isinstance(subset, str) else (list(subset) if subset else self.columns)
        schema = self._arrow_schema()
        from raydp_b200.sql import functions as F
        df = self
        for name in subset:
            t = schema.field(name).type
            numeric = pa.types.is_integer(t) or pa.types.is_floating(t)
            expr = None
            for old, new in mapping.items():
                if isinstance(old, str) != (pa.types.is_string(t) or pa.types.is_large_string(t)):
                    continue
                if isinstance(old, bool) != pa.types.is_boolean(t) or (not isinstance(old, (str, bool)) and not numeric):
                    continue
                cond = Column("col", (name,)) == old
                expr = F.when(cond, F.lit(new)) if expr is None else expr.when(cond, F.lit(new))
            if expr is not None:
                df = df.withColumn(name, expr.otherwise(Column("col", (name,))).cast(T.from_arrow_type(t)))
        return df

    def repartitionByRange(self, numPartitions, *cols):
        if not isinstance(numPartitions, int):
            cols = (numPartitions,) + cols
            numPartitions = self.rdd.getNumPartitions()
        from raydp_b200.sql.dataframe import _split_table
        from raydp_b200.sql.windowing import sort_table
        table = sort_table(self._collect_table(), cols)
        return self._session._from_tables(_split_table(table, numPartitions))

    # ------------------------------------------------------------------ statistics
    @property
    def stat(self):
        return _StatFunctions(self)

    def approxQuantile(self, col, probabilities, relativeError: float = 0.0):
        single = isinstance(col, str)
        names = [col] if single else list(col)
        table = self._collect_table()
        out = []
        for name in names:
            values = table.column(name).drop_null()
            if pa.types.is_floating(values.type):
                values = values.filter(pc.invert(pc.is_nan(values)))
            ordered = np.sort(values.to_numpy())
            if len(ordered) == 0:
                out.append([])
                continue
            ranks = [min(max(int(np.ceil(p * len(ordered))) - 1, 0), len(ordered) - 1) for p in probabilities]
            out.append([float(ordered[r]) for r in ranks])
        return out[0] if single else out

    def corr(self, col1: str, col2: str, method: Optional[str] = None) -> float:
        from raydp_b200.sql import functions as F
        return self.agg(F.corr(col1, col2).alias("c")).collect()[0][0]

    def cov(self, col1: str, col2: str) -> float:
        from raydp_b200.sql import functions as F
        return self.agg(F.covar_samp(col1, col2).alias("c")).collect()[0][0]

    def crosstab(self, col1: str, col2: str):
        from raydp_b200.sql import functions as F
        out = self.groupBy(col1).pivot(col2).agg(F.count("*"))
        renamed = out.withColumnRenamed(col1, f"{col1}_{col2}")
        return renamed.fillna(0)

    def freqItems(self, cols, support: float = 0.01):
        cols = [cols] if isinstance(cols, str) else list(cols)
        table = self._collect_table()
        n = max(table.num_rows, 1)
        data = {}
        for c in cols:
            counts = pc.value_counts(table.column(c).combine_chunks())
            items = [v for v, k in zip(pc.struct_field(counts, "values").to_pylist(),
                                       pc.struct_field(counts, "counts").to_pylist()) if k / n >= support]
            data[f"{c}_freqItems"] = pa.array([items], type=pa.list_(table.schema.field(c).type))
        return self._session._from_tables([pa.table(data)])

    def sampleBy(self, col, fractions: Dict[Any, float], seed: Optional[int] = None):
        name = col if isinstance(col, str) else col._output_name()
        table = self._collect_table()
        rng = np.random.default_rng(seed if seed is not None else 0x5eed)
        draw = rng.random(table.num_rows)
        keys = table.column(name).to_pylist()
        keep = np.array([d < fractions.get(k, 0.0) for d, k in zip(draw, keys)], dtype=bool)
        return self._session._from_tables([table.filter(pa.array(keep))])

    def summary(self, *statistics: str):
        stats = list(statistics) or ["count", "mean", "stddev", "min", "25%", "50%", "75%", "max"]
        table = self._collect_table()
        names = [f.name for f in table.schema if pa.types.is_integer(f.type) or pa.types.is_floating(f.type) or
                 pa.types.is_string(f.type)]
        rows: Dict[str, List[Optional[str]]] = {"summary": stats}
        for name in names:
            c = table.column(name).combine_chunks()
            numeric = not pa.types.is_string(c.type)
            live = c.drop_null()
            cells: List[Optional[str]] = []
            for st in stats:
                if st == "count":
                    cells.append(str(len(live)))
                elif st in ("min", "max"):
                    cells.append(None if len(live) == 0 else str(getattr(pc, st)(live).as_py()))
                elif not numeric or len(live) == 0:
                    cells.append(None)
                elif st == "mean":
                    cells.append(str(pc.mean(live).as_py()))
                elif st == "stddev":
                    cells.append(str(pc.stddev(live, ddof=1).as_py()) if len(live) > 1 else None)
                elif st.endswith("%"):
                    ordered = np.sort(live.to_numpy())
                    rank = min(max(int(np.ceil(float(st[:-1]) / 100.0 * len(ordered))) - 1, 0), len(ordered) - 1)
                    cells.append(str(ordered[rank].item()))
                else:
                    raise ValueError(f"{st} is not a recognised statistic")
            rows[name] = cells
        return self._session._from_tables([pa.table({k: pa.array(v, type=pa.string()) for k, v in rows.items()})])

    # ------------------------------------------------------------------ multi-level grouping
    def rollup(self, *cols):
        from raydp_b200.sql.dataframe import GroupingSets
        names = _names(cols)
        sets = [names[:i] for i in range(len(names), -1, -1)]
        return GroupingSets(self, names, sets)

    def cube(self, *cols):
        from raydp_b200.sql.dataframe import GroupingSets
        names = _names(cols)
        sets = [list(c) for r in range(len(names), -1, -1) for c in itertools.combinations(names, r)]
        return GroupingSets(self, names, sets)

    def writeTo(self, table: str):
        return _WriterV2(self, table)


def _chunks(items: list, n: int) -> List[list]:
    per, rem = divmod(len(items), n)
    out, off = [], 0
    for i in range(n):
        size = per + (1 if i < rem else 0)
        out.append(items[off:off + size])
        off += size
    return out


def _names(cols) -> List[str]:
    if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
        cols = tuple(cols[0])
    return [c if isinstance(c, str) else c._output_name() for c in cols]


class _StatFunctions:
    def __init__(self, df):
        self._df = df

    def approxQuantile(self, col, probabilities, relativeError=0.0):
        return self._df.approxQuantile(col, probabilities, relativeError)

    def corr(self, col1, col2, method=None):
        return self._df.corr(col1, col2, method)

    def cov(self, col1, col2):
        return self._df.cov(col1, col2)

    def crosstab(self, col1, col2):
        return self._df.crosstab(col1, col2)

    def freqItems(self, cols, support=0.01):
        return self._df.freqItems(cols, support)

    def sampleBy(self, col, fractions, seed=None):
        return self._df.sampleBy(col, fractions, seed)


class _WriterV2:
    """``df.writeTo("t").create() / append() / createOrReplace()`` over the session catalog's tables."""

    def __init__(self, df, table: str):
        self._df, self._table = df, table

    def using(self, provider: str):
        return self

    def option(self, key, value):
        return self

    def partitionedBy(self, *cols):
        return self

    def tableProperty(self, key, value):
        return self

    def create(self) -> None:
        self._df.write.saveAsTable(self._table, mode="errorifexists")

    def replace(self) -> None:
        self._df.write.saveAsTable(self._table, mode="overwrite")

    createOrReplace = replace

    def append(self) -> None:
        self._df.write.saveAsTable(self._table, mode="append")

    def overwritePartitions(self) -> None:
        self._df.write.saveAsTable(self._table, mode="overwrite")
