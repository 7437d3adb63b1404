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
            parts.append(T.table_to_rows(t))
        return RDD(self._df._session, parts)

    def mapPartitions(self, f, preservesPartitioning: bool = False) -> RDD:
        return self._as_rdd().mapPartitions(f)

    def map(self, f) -> RDD:
        return self._as_rdd().map(f)

    def collect(self):
        return self._df.collect()

    def count(self) -> int:
        return self._df.count()


# ---------------------------------------------------------------------------------------------------
# the wider RDD surface.  Narrow transformations stay lazy (chained partition functions, run on the
# executors); anything that needs all the data -- shuffles by key, sorts, set operations -- collects
# the partitions, works on them on the driver and re-slices, which is the right trade for the small
# RDDs DataFrame programs build on the side (lookup tables, id lists, parameter grids).
# ---------------------------------------------------------------------------------------------------
def _reslice(session, items: List[Any], n: int) -> "RDD":
    n = max(1, n)
    per, rem = divmod(len(items), n)
    parts, off = [], 0
    for i in range(n):
        size = per + (1 if i < rem else 0)
        parts.append(items[off:off + size])
        off += size
    return RDD(session, parts)


def _rdd_extras():
    import functools
    import itertools
    import operator
    import random

    def first(self):
        for part in self._collect_partitions():
            if part:
                return part[0]
        raise ValueError("RDD is empty")

    def take(self, num: int):
        return self.collect()[:num]

    def top(self, num: int, key=None):
        return sorted(self.collect(), key=key, reverse=True)[:num]

    def takeOrdered(self, num: int, key=None):
        return sorted(self.collect(), key=key)[:num]

    def takeSample(self, withReplacement: bool, num: int, seed=None):
        rng = random.Random(seed)
        items = self.collect()
        return [rng.choice(items) for _ in range(num)] if withReplacement else rng.sample(items, min(num, len(items)))

    def isEmpty(self) -> bool:
        return all(not p for p in self._collect_partitions())

    def reduce(self, f):
        items = self.collect()
        if not items:
            raise ValueError("Can not reduce() empty RDD")
        return functools.reduce(f, items)

    def fold(self, zeroValue, op):
        return functools.reduce(op, [functools.reduce(op, p, zeroValue) for p in self._collect_partitions()], zeroValue)

    def aggregate(self, zeroValue, seqOp, combOp):
        import copy
        partials = [functools.reduce(seqOp, p, copy.deepcopy(zeroValue)) for p in self._collect_partitions()]
        return functools.reduce(combOp, partials, copy.deepcopy(zeroValue))

    def sum_(self):
        return sum(self.collect())

    def mean(self):
        items = self.collect()
        return sum(items) / len(items)

    def max_(self, key=None):
        return max(self.collect(), key=key) if key else max(self.collect())

    def min_(self, key=None):
        return min(self.collect(), key=key) if key else min(self.collect())

    def stdev(self):
        import statistics
        return statistics.pstdev(self.collect())

    def variance(self):
        import statistics
        return statistics.pvariance(self.collect())

    def countByValue(self):
        out: dict = {}
        for x in self.collect():
            out[x] = out.get(x, 0) + 1
        return out

    def countByKey(self):
        out: dict = {}
        for k, _ in self.collect():
            out[k] = out.get(k, 0) + 1
        return out

    def collectAsMap(self):
        return dict(self.collect())

    def foreach(self, f) -> None:
        self.mapPartitions(lambda it, _f=f: [_f(x) for x in it] and []).collect()

    def foreachPartition(self, f) -> None:
        self.mapPartitions(lambda it, _f=f: _f(it) or []).collect()

    def glom(self):
        return self.mapPartitions(lambda it: [list(it)])

    def mapPartitionsWithIndex(self, f, preservesPartitioning: bool = False):
        parts = self._collect_partitions()
        return RDD(self._session, [list(f(i, iter(p))) for i, p in enumerate(parts)])

    def keys(self):
        return self.map(lambda kv: kv[0])

    def values(self):
        return self.map(lambda kv: kv[1])

    def keyBy(self, f):
        return self.map(lambda x, _f=f: (_f(x), x))

    def mapValues(self, f):
        return self.map(lambda kv, _f=f: (kv[0], _f(kv[1])))

    def flatMapValues(self, f):
        return self.flatMap(lambda kv, _f=f: ((kv[0], v) for v in _f(kv[1])))

    def zipWithIndex(self):
        return _reslice_like(self, list(zip(self.collect(), itertools.count())))

    def zipWithUniqueId(self):
        return self.zipWithIndex()

    def zip_(self, other):
        a, b = self.collect(), other.collect()
        if len(a) != len(b):
            raise ValueError("Can only zip RDDs with same number of elements in each partition")
        return _reslice_like(self, list(zip(a, b)))

    def _reslice_like(rdd, items):
        return _reslice(rdd._session, items, rdd.getNumPartitions())

    def distinct(self, numPartitions=None):
        seen, out = set(), []
        for x in self.collect():
            if x not in seen:
                seen.add(x)
                out.append(x)
        return _reslice(self._session, out, numPartitions or self.getNumPartitions())

    def union(self, other):
        return RDD(self._session, self._collect_partitions() + other._collect_partitions())

    def intersection(self, other):
        theirs = set(other.collect())
        return _reslice_like(self, [x for x in dict.fromkeys(self.collect()) if x in theirs])

    def subtract(self, other, numPartitions=None):
        theirs = set(other.collect())
        return _reslice_like(self, [x for x in self.collect() if x not in theirs])

    def cartesian(self, other):
        return _reslice_like(self, list(itertools.product(self.collect(), other.collect())))

    def sortBy(self, keyfunc, ascending: bool = True, numPartitions=None):
        return _reslice(self._session, sorted(self.collect(), key=keyfunc, reverse=not ascending),
                        numPartitions or self.getNumPartitions())

    def sortByKey(self, ascending: bool = True, numPartitions=None, keyfunc=lambda k: k):
        return sortBy(self, lambda kv: keyfunc(kv[0]), ascending, numPartitions)

    def groupBy(self, f, numPartitions=None):
        return groupByKey(keyBy(self, f), numPartitions)

    def groupByKey(self, numPartitions=None):
        groups: dict = {}
        for k, v in self.collect():
            groups.setdefault(k, []).append(v)
        return _reslice(self._session, list(groups.items()), numPartitions or self.getNumPartitions())

    def reduceByKey(self, func, numPartitions=None):
        acc: dict = {}
        for k, v in self.collect():
            acc[k] = func(acc[k], v) if k in acc else v
        return _reslice(self._session, list(acc.items()), numPartitions or self.getNumPartitions())

    def foldByKey(self, zeroValue, func, numPartitions=None):
        acc: dict = {}
        for k, v in self.collect():
            acc[k] = func(acc.get(k, zeroValue), v)
        return _reslice(self._session, list(acc.items()), numPartitions or self.getNumPartitions())

    def aggregateByKey(self, zeroValue, seqFunc, combFunc, numPartitions=None):
        import copy
        acc: dict = {}
        for k, v in self.collect():
            acc[k] = seqFunc(acc[k] if k in acc else copy.deepcopy(zeroValue), v)
        return _reslice(self._session, list(acc.items()), numPartitions or self.getNumPartitions())

    def combineByKey(self, createCombiner, mergeValue, mergeCombiners, numPartitions=None):
        acc: dict = {}
        for k, v in self.collect():
            acc[k] = mergeValue(acc[k], v) if k in acc else createCombiner(v)
        return _reslice(self._session, list(acc.items()), numPartitions or self.getNumPartitions())

    def _join(self, other, how: str):
        left: dict = {}
        for k, v in self.collect():
            left.setdefault(k, []).append(v)
        right: dict = {}
        for k, v in other.collect():
            right.setdefault(k, []).append(v)
        out = []
        for k in list(left) + [k for k in right if k not in left]:
            ls, rs = left.get(k), right.get(k)
            if ls and rs:
                out.extend((k, (a, b)) for a in ls for b in rs)
            elif ls and how in ("left", "full"):
                out.extend((k, (a, None)) for a in ls)
            elif rs and how in ("right", "full"):
                out.extend((k, (None, b)) for b in rs)
        return _reslice_like(self, out)

    def cogroup(self, other, numPartitions=None):
        left: dict = {}
        for k, v in self.collect():
            left.setdefault(k, []).append(v)
        right: dict = {}
        for k, v in other.collect():
            right.setdefault(k, []).append(v)
        keys = list(left) + [k for k in right if k not in left]
        return _reslice_like(self, [(k, (left.get(k, []), right.get(k, []))) for k in keys])

    def sample(self, withReplacement: bool, fraction: float, seed=None):
        rng = random.Random(seed)
        return _reslice_like(self, [x for x in self.collect() if rng.random() < fraction])

    def randomSplit(self, weights, seed=None):
        rng = random.Random(seed)
        total = float(sum(weights))
        bounds = list(itertools.accumulate(w / total for w in weights))
        buckets: List[List[Any]] = [[] for _ in weights]
        for x in self.collect():
            r = rng.random()
            buckets[next(i for i, b in enumerate(bounds) if r <= b or i == len(bounds) - 1)].append(x)
        return [_reslice_like(self, b) for b in buckets]

    def repartition(self, numPartitions: int):
        return _reslice(self._session, self.collect(), numPartitions)

    def coalesce(self, numPartitions: int, shuffle: bool = False):
        return _reslice(self._session, self.collect(), min(numPartitions, self.getNumPartitions()))

    def persist(self, storageLevel=None):
        if self._fns:                               # materialise once; later actions reuse the result
            self._partitions, self._fns = self._collect_partitions(), ()
        return self

    def unpersist(self, blocking: bool = False):
        return self

    def saveAsTextFile(self, path: str, compressionCodecClass=None) -> None:
        import os
        os.makedirs(path, exist_ok=False)
        for i, part in enumerate(self._collect_partitions()):
            with open(os.path.join(path, f"part-{i:05d}"), "w", encoding="utf-8") as fh:
                for x in part:
                    fh.write(str(x) + "\n")
        open(os.path.join(path, "_SUCCESS"), "w").close()

    def saveAsPickleFile(self, path: str, batchSize: int = 10) -> None:
        import os
        import pickle
        os.makedirs(path, exist_ok=False)
        for i, part in enumerate(self._collect_partitions()):
            with open(os.path.join(path, f"part-{i:05d}"), "wb") as fh:
                pickle.dump(part, fh)

    def toLocalIterator(self, prefetchPartitions: bool = False):
        return iter(self.collect())

    def histogram(self, buckets):
        import numpy as np
        counts, edges = np.histogram(np.asarray(self.collect(), dtype=float), bins=buckets)
        return [e.item() for e in edges], [int(c) for c in counts]

    def stats(self):
        import statistics
        items = self.collect()
        return {"count": len(items), "mean": statistics.fmean(items), "stdev": statistics.pstdev(items),
                "max": max(items), "min": min(items)}

    methods = dict(
        first=first, take=take, top=top, takeOrdered=takeOrdered, takeSample=takeSample, isEmpty=isEmpty,
        reduce=reduce, fold=fold, aggregate=aggregate, sum=sum_, mean=mean, max=max_, min=min_, stdev=stdev,
        variance=variance, countByValue=countByValue, countByKey=countByKey, collectAsMap=collectAsMap,
        foreach=foreach, foreachPartition=foreachPartition, glom=glom, mapPartitionsWithIndex=mapPartitionsWithIndex,
        keys=keys, values=values, keyBy=keyBy, mapValues=mapValues, flatMapValues=flatMapValues,
        zipWithIndex=zipWithIndex, zipWithUniqueId=zipWithUniqueId, zip=zip_, distinct=distinct, union=union,
        intersection=intersection, subtract=subtract, cartesian=cartesian, sortBy=sortBy, sortByKey=sortByKey,
        groupBy=groupBy, groupByKey=groupByKey, reduceByKey=reduceByKey, foldByKey=foldByKey,
        aggregateByKey=aggregateByKey, combineByKey=combineByKey, cogroup=cogroup, sample=sample,
        randomSplit=randomSplit, repartition=repartition, coalesce=coalesce, persist=persist, cache=persist,
        unpersist=unpersist, saveAsTextFile=saveAsTextFile, saveAsPickleFile=saveAsPickleFile,
        toLocalIterator=toLocalIterator, histogram=histogram, stats=stats,
        join=lambda self, other, numPartitions=None: _join(self, other, "inner"),
        leftOuterJoin=lambda self, other, numPartitions=None: _join(self, other, "left"),
        rightOuterJoin=lambda self, other, numPartitions=None: _join(self, other, "right"),
        fullOuterJoin=lambda self, other, numPartitions=None: _join(self, other, "full"),
        __add__=union, context=property(lambda self: self._session.sparkContext),
        id=lambda self: id(self), name=lambda self: None, setName=lambda self, name: self,
        getStorageLevel=lambda self: None, is_cached=property(lambda self: not self._fns),
    )
    for name, fn in methods.items():
        if name not in RDD.__dict__:
            setattr(RDD, name, fn)

    # df.rdd: everything the plain RDD offers, over the rows
    def delegate(name):
        def call(self, *args, **kwargs):
            return getattr(self._as_rdd(), name)(*args, **kwargs)
        call.__name__ = name
        return call
    for name, fn in methods.items():
        if name not in DataFrameRDD.__dict__ and callable(fn) and not name.startswith("__"):
            setattr(DataFrameRDD, name, delegate(name))
    for name in ("filter", "flatMap", "toDF"):
        if name not in DataFrameRDD.__dict__:
            setattr(DataFrameRDD, name, delegate(name))


_rdd_extras()
