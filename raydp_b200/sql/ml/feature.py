"""``pyspark.ml.feature`` stages on the built-in engine.

Vector columns are ``array<double>``.  Row-local stages become one Arrow kernel per partition (an
``arrowfn`` column expression, so they fuse into the lazy plan and run on the executors); stages that
need statistics (``StringIndexer``, scalers, ``Imputer``, ``QuantileDiscretizer``) compute them with one
aggregation in ``fit``.
"""
from __future__ import annotations

import re
import zlib
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

from raydp_b200.sql.column import Column
from raydp_b200.sql.ml.base import Estimator, Model, Transformer


def _col(name: str) -> Column:
    return Column("col", (name,))


def _arrow_expr(fn: Callable, *cols: str, name: str = "fn", empty: Optional[pa.DataType] = None) -> Column:
    """``fn(*arrow arrays) -> arrow array`` as a column expression.  Schema inference evaluates plans on
    zero rows, where a vector has no width yet: ``empty`` is the result type reported for that case."""
    empty_type = empty if empty is not None else pa.list_(pa.float64())

    def guarded(*arrays, _fn=fn, _empty=empty_type):
        if arrays and len(arrays[0]) == 0:
            return pa.array([], type=_empty)
        return _fn(*arrays)
    guarded.__name__ = name
    return Column("arrowfn", (guarded, tuple(_col(c) for c in cols)))


def _to_matrix(arr: pa.Array) -> np.ndarray:
    """``array<double>`` column -> (rows, width) float64 matrix (rows must share a width; nulls -> NaN)."""
    if pa.types.is_list(arr.type) or pa.types.is_large_list(arr.type) or pa.types.is_fixed_size_list(arr.type):
        n = len(arr)
        if n == 0:
            return np.zeros((0, 0))
        lengths = np.asarray(pc.fill_null(pc.list_value_length(arr), 0))
        width = int(lengths.max())
        flat = pc.cast(pc.list_flatten(arr), pa.float64()).to_numpy(zero_copy_only=False)
        if (lengths == width).all():
            return np.asarray(flat, dtype=np.float64).reshape(n, width)
        out = np.full((n, width), np.nan)
        pos = 0
        for i, ln in enumerate(lengths):
            out[i, :ln] = flat[pos:pos + ln]
            pos += ln
        return out
    return np.asarray(pc.cast(arr, pa.float64()).to_numpy(zero_copy_only=False), dtype=np.float64).reshape(-1, 1)


def _from_matrix(m: np.ndarray) -> pa.Array:
    n, width = m.shape
    offsets = pa.array(np.arange(0, (n + 1) * width, width, dtype=np.int32))
    return pa.ListArray.from_arrays(offsets, pa.array(np.ascontiguousarray(m, dtype=np.float64).reshape(-1)))


def _inputs(stage, single: str = "inputCol", multi: str = "inputCols") -> List[str]:
    if stage.hasParam(multi) and stage.isSet(multi):
        return list(stage.getOrDefault(multi))
    return [stage.getOrDefault(single)]


def _outputs(stage, ins: List[str], single: str = "outputCol", multi: str = "outputCols") -> List[str]:
    if stage.hasParam(multi) and stage.isSet(multi):
        return list(stage.getOrDefault(multi))
    out = stage.getOrDefault(single)
    if out is None:
        raise ValueError(f"{type(stage).__name__}: {single} is not set")
    return [out] if len(ins) == 1 else [f"{out}_{c}" for c in ins]


# ------------------------------------------------------------------------------------------ assembling
class VectorAssembler(Transformer):
    _defaults = {"inputCols": None, "outputCol": None, "handleInvalid": "error"}

    def __init__(self, *, inputCols=None, outputCol=None, handleInvalid=None):
        super().__init__(inputCols=inputCols, outputCol=outputCol, handleInvalid=handleInvalid)

    def _transform(self, dataset):
        cols = list(self.getOrDefault("inputCols"))
        invalid = self.getOrDefault("handleInvalid")

        def assemble(*arrays):
            parts = [_to_matrix(a) for a in arrays]
            m = np.concatenate(parts, axis=1) if parts else np.zeros((0, 0))
            bad = np.isnan(m).any(axis=1) if m.size else np.zeros(len(m), dtype=bool)
            if bad.any() and invalid == "error":
                raise ValueError("VectorAssembler: null / NaN in the input columns; use handleInvalid='skip' or 'keep'")
            out = _from_matrix(m)
            if bad.any() and invalid == "skip":
                out = pc.if_else(pa.array(bad), pa.nulls(len(out), out.type), out)
            return out
        out = dataset.withColumn(self.getOrDefault("outputCol"), _arrow_expr(assemble, *cols, name="assemble"))
        if invalid == "skip":
            out = out.filter(_col(self.getOrDefault("outputCol")).isNotNull())
        return out


class VectorSlicer(Transformer):
    _defaults = {"inputCol": None, "outputCol": None, "indices": None}

    def __init__(self, *, inputCol=None, outputCol=None, indices=None, names=None):
        super().__init__(inputCol=inputCol, outputCol=outputCol, indices=indices)

    def _transform(self, dataset):
        idx = list(self.getOrDefault("indices"))
        return dataset.withColumn(self.getOrDefault("outputCol"), _arrow_expr(
            lambda a: _from_matrix(_to_matrix(a)[:, idx]), self.getOrDefault("inputCol"), name="slice"))


# ------------------------------------------------------------------------------------------ categories
class StringIndexer(Estimator):
    _defaults = {"inputCol": None, "outputCol": None, "inputCols": None, "outputCols": None,
                 "handleInvalid": "error", "stringOrderType": "frequencyDesc"}

    def __init__(self, *, inputCol=None, outputCol=None, inputCols=None, outputCols=None, handleInvalid=None,
                 stringOrderType=None):
        super().__init__(inputCol=inputCol, outputCol=outputCol, inputCols=inputCols, outputCols=outputCols,
                         handleInvalid=handleInvalid, stringOrderType=stringOrderType)

    def _fit(self, dataset) -> "StringIndexerModel":
        from raydp_b200.sql import functions as F
        ins = _inputs(self)
        order = self.getOrDefault("stringOrderType")
        labels = []
        for c in ins:
            counts = dataset.filter(_col(c).isNotNull()).groupBy(c).agg(F.count("*").alias("n")).collect()
            pairs = [(str(r[0]), r[1]) for r in counts]
            if order == "frequencyDesc":
                pairs.sort(key=lambda p: (-p[1], p[0]))
            elif order == "frequencyAsc":
                pairs.sort(key=lambda p: (p[1], p[0]))
            elif order == "alphabetDesc":
                pairs.sort(key=lambda p: p[0], reverse=True)
            else:
                pairs.sort(key=lambda p: p[0])
            labels.append([p[0] for p in pairs])
        model = StringIndexerModel(labels)
        return model._set(**self._values)


class StringIndexerModel(Model):
    _defaults = dict(StringIndexer._defaults)

    def __init__(self, labels_array: List[List[str]]):
        super().__init__()
        self.labelsArray = [list(ls) for ls in labels_array]

    @property
    def labels(self) -> List[str]:
        return self.labelsArray[0]

    @classmethod
    def from_labels(cls, labels, inputCol, outputCol=None, handleInvalid=None):
        return cls([list(labels)])._set(inputCol=inputCol, outputCol=outputCol, handleInvalid=handleInvalid)

    def _transform(self, dataset):
        ins = _inputs(self)
        outs = _outputs(self, ins)
        invalid = self.getOrDefault("handleInvalid")
        for c, out, labels in zip(ins, outs, self.labelsArray):
            value_set = pa.array(labels, type=pa.string())
            unseen = float(len(labels))

            def index(a, _vs=value_set, _unseen=unseen, _c=c):
                pos = pc.index_in(pc.cast(a, pa.string()), value_set=_vs)
                missing = np.asarray(pc.is_null(pos))
                if missing.any():
                    if invalid == "error":
                        raise ValueError(f"StringIndexer: unseen label or null in column {_c!r}; "
                                         "set handleInvalid to 'skip' or 'keep'")
                    if invalid == "keep":
                        pos = pc.fill_null(pos, int(_unseen))
                return pc.cast(pos, pa.float64())
            dataset = dataset.withColumn(out, _arrow_expr(index, c, name="string_index", empty=pa.float64()))
            if invalid == "skip":
                dataset = dataset.filter(_col(out).isNotNull())
        return dataset


class IndexToString(Transformer):
    _defaults = {"inputCol": None, "outputCol": None, "labels": None}

    def __init__(self, *, inputCol=None, outputCol=None, labels=None):
        super().__init__(inputCol=inputCol, outputCol=outputCol, labels=labels)

    def _transform(self, dataset):
        labels = pa.array(list(self.getOrDefault("labels")), type=pa.string())
        return dataset.withColumn(self.getOrDefault("outputCol"), _arrow_expr(
            lambda a: labels.take(pc.cast(a, pa.int64())), self.getOrDefault("inputCol"), name="index_to_string", empty=pa.string()))


class OneHotEncoder(Estimator):
    _defaults = {"inputCol": None, "outputCol": None, "inputCols": None, "outputCols": None,
                 "dropLast": True, "handleInvalid": "error"}

    def __init__(self, *, inputCol=None, outputCol=None, inputCols=None, outputCols=None, dropLast=None,
                 handleInvalid=None):
        super().__init__(inputCol=inputCol, outputCol=outputCol, inputCols=inputCols, outputCols=outputCols,
                         dropLast=dropLast, handleInvalid=handleInvalid)

    def _fit(self, dataset) -> "OneHotEncoderModel":
        from raydp_b200.sql import functions as F
        ins = _inputs(self)
        row = dataset.agg(*[F.max(c).alias(c) for c in ins]).collect()[0]
        sizes = [int(row[c]) + 1 if row[c] is not None else 0 for c in ins]
        return OneHotEncoderModel(sizes)._set(**self._values)


class OneHotEncoderModel(Model):
    _defaults = dict(OneHotEncoder._defaults)

    def __init__(self, category_sizes: List[int]):
        super().__init__()
        self.categorySizes = list(category_sizes)

    def _transform(self, dataset):
        ins = _inputs(self)
        outs = _outputs(self, ins)
        drop_last = bool(self.getOrDefault("dropLast"))
        keep_invalid = self.getOrDefault("handleInvalid") == "keep"
        for c, out, size in zip(ins, outs, self.categorySizes):
            width = size + (1 if keep_invalid else 0) - (1 if drop_last else 0)

            def encode(a, _size=size, _width=width, _c=c):
                idx = np.asarray(pc.fill_null(pc.cast(a, pa.int64()), -1))
                bad = (idx < 0) | (idx >= _size)
                if bad.any() and not keep_invalid:
                    raise ValueError(f"OneHotEncoder: category index outside [0, {_size}) in column {_c!r}")
                idx = np.where(bad, _size, idx)
                m = np.zeros((len(idx), max(_width, 0)))
                inside = idx < _width
                m[np.flatnonzero(inside), idx[inside]] = 1.0
                return _from_matrix(m)
            dataset = dataset.withColumn(out, _arrow_expr(encode, c, name="one_hot"))
        return dataset


# ------------------------------------------------------------------------------------------ scaling
class _VectorStatsEstimator(Estimator):
    """Fits per-dimension statistics of an ``array<double>`` column in one pass over the partitions."""

    _defaults = {"inputCol": None, "outputCol": None}

    def _stats(self, dataset) -> Dict[str, np.ndarray]:
        col = self.getOrDefault("inputCol")

        def per_partition(rows):
            import numpy as _np
            vals = [r[0] for r in rows if r[0] is not None]
            if not vals:
                return []
            m = _np.asarray(vals, dtype=_np.float64)
            return [(len(m), m.sum(axis=0), (m * m).sum(axis=0), m.min(axis=0), m.max(axis=0), _np.abs(m).max(axis=0))]
        parts = dataset.select(col).rdd.mapPartitions(per_partition).collect()
        if not parts:
            raise ValueError(f"{type(self).__name__}: column {col!r} has no rows to fit on")
        n = sum(p[0] for p in parts)
        s = np.sum([p[1] for p in parts], axis=0)
        sq = np.sum([p[2] for p in parts], axis=0)
        mean = s / n
        var = np.maximum(sq - n * mean * mean, 0.0) / max(n - 1, 1)
        return {"n": n, "mean": mean, "std": np.sqrt(var), "min": np.min([p[3] for p in parts], axis=0),
                "max": np.max([p[4] for p in parts], axis=0), "maxabs": np.max([p[5] for p in parts], axis=0)}


class _VectorMapModel(Model):
    _defaults = {"inputCol": None, "outputCol": None}

    def _map(self, m: np.ndarray) -> np.ndarray:
        raise NotImplementedError

    def _transform(self, dataset):
        return dataset.withColumn(self.getOrDefault("outputCol"), _arrow_expr(
            lambda a: _from_matrix(self._map(_to_matrix(a))), self.getOrDefault("inputCol"),
            name=type(self).__name__))


class StandardScaler(_VectorStatsEstimator):
    _defaults = {"withMean": False, "withStd": True}

    def __init__(self, *, withMean=None, withStd=None, inputCol=None, outputCol=None):
        super().__init__(withMean=withMean, withStd=withStd, inputCol=inputCol, outputCol=outputCol)

    def _fit(self, dataset) -> "StandardScalerModel":
        st = self._stats(dataset)
        return StandardScalerModel(st["mean"], st["std"])._set(**self._values)


class StandardScalerModel(_VectorMapModel):
    _defaults = {"withMean": False, "withStd": True}

    def __init__(self, mean: np.ndarray, std: np.ndarray):
        super().__init__()
        self.mean, self.std = np.asarray(mean), np.asarray(std)

    def _map(self, m):
        if self.getOrDefault("withMean"):
            m = m - self.mean
        if self.getOrDefault("withStd"):
            m = m * np.where(self.std > 0, 1.0 / np.where(self.std > 0, self.std, 1.0), 0.0)
        return m


class MinMaxScaler(_VectorStatsEstimator):
    _defaults = {"min": 0.0, "max": 1.0}

    def __init__(self, *, min=None, max=None, inputCol=None, outputCol=None):  # noqa: A002
        super().__init__(min=min, max=max, inputCol=inputCol, outputCol=outputCol)

    def _fit(self, dataset) -> "MinMaxScalerModel":
        st = self._stats(dataset)
        return MinMaxScalerModel(st["min"], st["max"])._set(**self._values)


class MinMaxScalerModel(_VectorMapModel):
    _defaults = {"min": 0.0, "max": 1.0}

    def __init__(self, original_min: np.ndarray, original_max: np.ndarray):
        super().__init__()
        self.originalMin, self.originalMax = np.asarray(original_min), np.asarray(original_max)

    def _map(self, m):
        lo, hi = float(self.getOrDefault("min")), float(self.getOrDefault("max"))
        span = self.originalMax - self.originalMin
        unit = np.where(span > 0, (m - self.originalMin) / np.where(span > 0, span, 1.0), 0.5)
        return unit * (hi - lo) + lo


class MaxAbsScaler(_VectorStatsEstimator):
    def __init__(self, *, inputCol=None, outputCol=None):
        super().__init__(inputCol=inputCol, outputCol=outputCol)

    def _fit(self, dataset) -> "MaxAbsScalerModel":
        return MaxAbsScalerModel(self._stats(dataset)["maxabs"])._set(**self._values)


class MaxAbsScalerModel(_VectorMapModel):
    def __init__(self, max_abs: np.ndarray):
        super().__init__()
        self.maxAbs = np.asarray(max_abs)

    def _map(self, m):
        return m / np.where(self.maxAbs > 0, self.maxAbs, 1.0)


class Normalizer(Transformer):
    _defaults = {"inputCol": None, "outputCol": None, "p": 2.0}

    def __init__(self, *, p=None, inputCol=None, outputCol=None):
        super().__init__(p=p, inputCol=inputCol, outputCol=outputCol)

    def _transform(self, dataset):
        p = float(self.getOrDefault("p"))

        def normalise(a):
            m = _to_matrix(a)
            norms = np.linalg.norm(m, ord=p, axis=1, keepdims=True) if m.size else np.ones((len(m), 1))
            return _from_matrix(m / np.where(norms > 0, norms, 1.0))
        return dataset.withColumn(self.getOrDefault("outputCol"),
                                  _arrow_expr(normalise, self.getOrDefault("inputCol"), name="normalize"))


# ------------------------------------------------------------------------------------------ numeric columns
class Imputer(Estimator):
    _defaults = {"inputCol": None, "outputCol": None, "inputCols": None, "outputCols": None, "strategy": "mean",
                 "missingValue": float("nan")}

    def __init__(self, *, strategy=None, missingValue=None, inputCols=None, outputCols=None, inputCol=None,
                 outputCol=None):
        super().__init__(strategy=strategy, missingValue=missingValue, inputCols=inputCols, outputCols=outputCols,
                         inputCol=inputCol, outputCol=outputCol)

    def _fit(self, dataset) -> "ImputerModel":
        from raydp_b200.sql import functions as F
        ins = _inputs(self)
        strategy = self.getOrDefault("strategy")
        missing = self.getOrDefault("missingValue")
        fills = []
        for c in ins:
            live = dataset.filter(_col(c).isNotNull() & ~F.isnan(_col(c).cast("double")))
            if missing == missing:                       # a non-NaN sentinel
                live = live.filter(_col(c) != missing)
            agg = {"mean": F.avg(c), "median": F.percentile_approx(c, 0.5), "mode": F.mode(c)}[strategy]
            fills.append(live.agg(agg.alias("v")).collect()[0][0])
        return ImputerModel(dict(zip(ins, fills)))._set(**self._values)


class ImputerModel(Model):
    _defaults = dict(Imputer._defaults)

    def __init__(self, surrogates: Dict[str, Any]):
        super().__init__()
        self.surrogates = dict(surrogates)

    def _transform(self, dataset):
        from raydp_b200.sql import functions as F
        ins = _inputs(self)
        outs = _outputs(self, ins)
        missing = self.getOrDefault("missingValue")
        for c, out in zip(ins, outs):
            fill = self.surrogates[c]
            dtype = dict(dataset.dtypes)[c]
            value = _col(c).cast("double")
            hole = _col(c).isNull() | F.isnan(value)
            if missing == missing:
                hole = hole | (value == float(missing))
            dataset = dataset.withColumn(out, F.when(hole, F.lit(fill)).otherwise(_col(c)).cast(dtype))
        return dataset


class Bucketizer(Transformer):
    _defaults = {"inputCol": None, "outputCol": None, "splits": None, "handleInvalid": "error"}

    def __init__(self, *, splits=None, inputCol=None, outputCol=None, handleInvalid=None):
        super().__init__(splits=splits, inputCol=inputCol, outputCol=outputCol, handleInvalid=handleInvalid)

    def _transform(self, dataset):
        splits = np.asarray(self.getOrDefault("splits"), dtype=np.float64)
        invalid = self.getOrDefault("handleInvalid")

        def bucket(a):
            x = pc.cast(a, pa.float64()).to_numpy(zero_copy_only=False)
            idx = np.searchsorted(splits, x, side="right") - 1
            idx = np.where(x == splits[-1], len(splits) - 2, idx)         # the last bucket includes its upper bound
            bad = np.isnan(x) | (idx < 0) | (idx > len(splits) - 2)
            if bad.any():
                if invalid == "error":
                    raise ValueError("Bucketizer: value outside the splits (or NaN); set handleInvalid")
                if invalid == "keep":
                    idx = np.where(bad, len(splits) - 1, idx)
                    bad = np.zeros(len(x), dtype=bool)
            return pa.array(idx.astype(np.float64), mask=bad)
        out = dataset.withColumn(self.getOrDefault("outputCol"),
                                 _arrow_expr(bucket, self.getOrDefault("inputCol"), name="bucketize", empty=pa.float64()))
        return out.filter(_col(self.getOrDefault("outputCol")).isNotNull()) if invalid == "skip" else out


class QuantileDiscretizer(Estimator):
    _defaults = {"inputCol": None, "outputCol": None, "numBuckets": 2, "relativeError": 0.001, "handleInvalid": "error"}

    def __init__(self, *, numBuckets=None, inputCol=None, outputCol=None, relativeError=None, handleInvalid=None):
        super().__init__(numBuckets=numBuckets, inputCol=inputCol, outputCol=outputCol, relativeError=relativeError,
                         handleInvalid=handleInvalid)

    def _fit(self, dataset) -> Bucketizer:
        k = int(self.getOrDefault("numBuckets"))
        cuts = dataset.approxQuantile(self.getOrDefault("inputCol"), [i / k for i in range(1, k)], 0.0)
        splits = [-float("inf")] + sorted(set(cuts)) + [float("inf")]
        return Bucketizer(splits=splits, inputCol=self.getOrDefault("inputCol"), outputCol=self.getOrDefault("outputCol"),
                          handleInvalid=self.getOrDefault("handleInvalid"))


class Binarizer(Transformer):
    _defaults = {"inputCol": None, "outputCol": None, "threshold": 0.0}

    def __init__(self, *, threshold=None, inputCol=None, outputCol=None):
        super().__init__(threshold=threshold, inputCol=inputCol, outputCol=outputCol)

    def _transform(self, dataset):
        from raydp_b200.sql import functions as F
        t = float(self.getOrDefault("threshold"))
        c = self.getOrDefault("inputCol")
        return dataset.withColumn(self.getOrDefault("outputCol"), F.when(_col(c) > t, 1.0).otherwise(0.0))


# ------------------------------------------------------------------------------------------ text
class Tokenizer(Transformer):
    _defaults = {"inputCol": None, "outputCol": None}

    def __init__(self, *, inputCol=None, outputCol=None):
        super().__init__(inputCol=inputCol, outputCol=outputCol)

    def _transform(self, dataset):
        from raydp_b200.sql import functions as F
        return dataset.withColumn(self.getOrDefault("outputCol"),
                                  F.split(F.lower(F.trim(_col(self.getOrDefault("inputCol")))), r"\s+"))


class RegexTokenizer(Transformer):
    _defaults = {"inputCol": None, "outputCol": None, "pattern": r"\s+", "gaps": True, "minTokenLength": 1,
                 "toLowercase": True}

    def __init__(self, *, minTokenLength=None, gaps=None, pattern=None, inputCol=None, outputCol=None, toLowercase=None):
        super().__init__(minTokenLength=minTokenLength, gaps=gaps, pattern=pattern, inputCol=inputCol,
                         outputCol=outputCol, toLowercase=toLowercase)

    def _transform(self, dataset):
        rx = re.compile(self.getOrDefault("pattern"))
        gaps, lower = bool(self.getOrDefault("gaps")), bool(self.getOrDefault("toLowercase"))
        min_len = int(self.getOrDefault("minTokenLength"))

        def tokens(a):
            out = []
            for s in a.to_pylist():
                if s is None:
                    out.append(None)
                    continue
                s = s.lower() if lower else s
                found = rx.split(s) if gaps else rx.findall(s)
                out.append([t for t in found if len(t) >= min_len])
            return pa.array(out, type=pa.list_(pa.string()))
        return dataset.withColumn(self.getOrDefault("outputCol"),
                                  _arrow_expr(tokens, self.getOrDefault("inputCol"), name="regex_tokenize",
                                              empty=pa.list_(pa.string())))


class StopWordsRemover(Transformer):
    _defaults = {"inputCol": None, "outputCol": None, "stopWords": None, "caseSensitive": False}
    _ENGLISH = ("a an and are as at be but by for if in into is it no not of on or such that the their then there "
                "these they this to was will with i me my we our you your he she him her its").split()

    def __init__(self, *, inputCol=None, outputCol=None, stopWords=None, caseSensitive=None):
        super().__init__(inputCol=inputCol, outputCol=outputCol, stopWords=stopWords, caseSensitive=caseSensitive)

    @staticmethod
    def loadDefaultStopWords(language: str = "english") -> List[str]:
        return list(StopWordsRemover._ENGLISH)

    def _transform(self, dataset):
        case = bool(self.getOrDefault("caseSensitive"))
        words = self.getOrDefault("stopWords") or self._ENGLISH
        stop = set(words if case else [w.lower() for w in words])

        def strip(a):
            return pa.array([None if row is None else [t for t in row if (t if case else t.lower()) not in stop]
                             for row in a.to_pylist()], type=pa.list_(pa.string()))
        return dataset.withColumn(self.getOrDefault("outputCol"),
                                  _arrow_expr(strip, self.getOrDefault("inputCol"), name="remove_stop_words",
                                              empty=pa.list_(pa.string())))


class HashingTF(Transformer):
    _defaults = {"inputCol": None, "outputCol": None, "numFeatures": 1 << 18, "binary": False}

    def __init__(self, *, numFeatures=None, binary=None, inputCol=None, outputCol=None):
        super().__init__(numFeatures=numFeatures, binary=binary, inputCol=inputCol, outputCol=outputCol)

    def indexOf(self, term) -> int:
        return zlib.crc32(str(term).encode()) % int(self.getOrDefault("numFeatures"))

    def _transform(self, dataset):
        width, binary = int(self.getOrDefault("numFeatures")), bool(self.getOrDefault("binary"))

        def tf(a):
            rows = a.to_pylist()
            m = np.zeros((len(rows), width))
            for i, row in enumerate(rows):
                for term in row or []:
                    j = zlib.crc32(str(term).encode()) % width
                    m[i, j] = 1.0 if binary else m[i, j] + 1.0
            return _from_matrix(m)
        return dataset.withColumn(self.getOrDefault("outputCol"),
                                  _arrow_expr(tf, self.getOrDefault("inputCol"), name="hashing_tf"))


class SQLTransformer(Transformer):
    _defaults = {"statement": None}

    def __init__(self, *, statement=None):
        super().__init__(statement=statement)

    def _transform(self, dataset):
        import uuid
        view = f"__sqlt_{uuid.uuid4().hex[:8]}"
        dataset.createOrReplaceTempView(view)
        try:
            return dataset.sparkSession.sql(self.getOrDefault("statement").replace("__THIS__", view))
        finally:
            dataset.sparkSession.catalog.dropTempView(view)
