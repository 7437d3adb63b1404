"""Minimal RDD: just enough for the reference's usages.

Parity: ``df.rdd.getNumPartitions()`` (reference ``spark/dataset.py:177``),
``sparkContext.parallelize(list[, n])``, ``rdd.map/mapPartitions(f).collect()`` and
``createDataFrame(rdd, names)`` (reference ``test_from_spark.py:24-25``,
``examples/test_pyfiles_main.py:45-53``).
"""
from __future__ import annotations

from typing import Any, Callable, Iterable, List

import pyarrow as pa

from raydp_b200.sql import types as T
from raydp_b200.sql.plan import FuncSource, Plan


class RDD:
    def __init__(self, session, partitions: List[List[Any]], fns: tuple = ()):
        self._session = session
        self._partitions = partitions
        self._fns = fns  # chain of partition-level callables: iterator -> iterable

    def getNumPartitions(self) -> int:
        return len(self._partitions)

    def mapPartitions(self, f: Callable[[Iterable], Iterable], preservesPartitioning: bool = False) -> "RDD":
        return RDD(self._session, self._partitions, self._fns + (f,))

    def map(self, f: Callable[[Any], Any]) -> "RDD":
        return self.mapPartitions(lambda it, _f=f: (_f(x) for x in it))

    def filter(self, f: Callable[[Any], bool]) -> "RDD":
        return self.mapPartitions(lambda it, _f=f: (x for x in it if _f(x)))

    def flatMap(self, f: Callable[[Any], Iterable]) -> "RDD":
        return self.mapPartitions(lambda it, _f=f: (y for x in it for y in _f(x)))

    def _collect_partitions(self) -> List[List[Any]]:
        if not self._fns:
            return [list(p) for p in self._partitions]
        parts, fns = self._partitions, self._fns
        schema = pa.schema([pa.field("blob", pa.binary())])

        def load(i, _parts=parts):
            import cloudpickle
            return pa.table({"blob": [cloudpickle.dumps(_parts[i])]})

        def run(table, _fns=fns):
            import cloudpickle
            data = iter(cloudpickle.loads(table.column(0)[0].as_py()))
            for fn in _fns:
                data = iter(fn(data))
            return list(data)
        plan = Plan(FuncSource(load, len(parts), schema))
        return self._session._backend.run(plan, "py_partition", {"fn": run})

    def collect(self) -> List[Any]:
        return [x for part in self._collect_partitions() for x in part]

    def count(self) -> int:
        return sum(len(p) for p in self._collect_partitions())

    def toDF(self, schema=None):
        return self._session.createDataFrame(self, schema)


class DataFrameRDD:
    """``df.rdd`` view: rows of a DataFrame as an RDD of :class:`Row`."""

    def __init__(self, df):
        self._df = df

    def getNumPartitions(self) -> int:
        self._df._materialise_cache()
        return self._df._plan.num_partitions

    def _as_rdd(self) -> RDD:
        tables = self._df._run("collect")
        parts = []
        for t in tables:
            names = t.column_names
            cols = [c.to_pylist() for c in t.columns]
            parts.append([T.Row._make(names, vals) for vals in zip(*cols)] if cols else [])
        return RDD(self._df._session, parts)

    def mapPartitions(self, f, preservesPartitioning: bool = False) -> RDD:
        return self._as_rdd().mapPartitions(f)

    def map(self, f) -> RDD:
        return self._as_rdd().map(f)

    def collect(self):
        return self._df.collect()

    def count(self) -> int:
        return self._df.count()
