"""pandas-on-Spark subset (``pyspark.pandas`` stand-in).

Parity: exactly what the reference's tests do with ``pyspark.pandas`` -- ``ps.range``, column
assignment / arithmetic, ``astype("float")``, ``.to_spark()``, ``ps.DataFrame(sdf)``
(reference ``test_torch.py:40-46``, ``test_spark_utils.py:30-72``, ``utils.py:82-83,122``).
Columns are lazy expressions over the underlying engine DataFrame.
"""
from __future__ import annotations

from typing import List, Optional

from raydp_b200.sql.column import Column
from raydp_b200.sql.dataframe import DataFrame as _SparkDF


class Series:
    def __init__(self, expr: Column, name: Optional[str] = None):
        self._expr = expr
        self.name = name

    def _wrap(self, expr: Column) -> "Series":
        return Series(expr, self.name)

    @staticmethod
    def _unwrap(other):
        return other._expr if isinstance(other, Series) else other

    def __add__(self, o): return self._wrap(self._expr + self._unwrap(o))
    def __radd__(self, o): return self._wrap(self._unwrap(o) + self._expr)
    def __sub__(self, o): return self._wrap(self._expr - self._unwrap(o))
    def __rsub__(self, o): return self._wrap(self._unwrap(o) - self._expr)
    def __mul__(self, o): return self._wrap(self._expr * self._unwrap(o))
    def __rmul__(self, o): return self._wrap(self._unwrap(o) * self._expr)
    def __truediv__(self, o): return self._wrap(self._expr / self._unwrap(o))
    def __rtruediv__(self, o): return self._wrap(self._unwrap(o) / self._expr)
    def __lt__(self, o): return self._wrap(self._expr < self._unwrap(o))
    def __le__(self, o): return self._wrap(self._expr <= self._unwrap(o))
    def __gt__(self, o): return self._wrap(self._expr > self._unwrap(o))
    def __ge__(self, o): return self._wrap(self._expr >= self._unwrap(o))

    def astype(self, dtype) -> "Series":
        name = getattr(dtype, "__name__", str(dtype)).lower()
        mapping = {"float": "double", "float64": "double", "float32": "float", "int": "bigint",
                   "int64": "bigint", "int32": "int", "str": "string", "bool": "boolean"}
        return self._wrap(self._expr.cast(mapping.get(name, name)))


class DataFrame:
    def __init__(self, data=None, columns: Optional[List[str]] = None):
        if isinstance(data, _SparkDF):
            self._sdf = data
        elif isinstance(data, DataFrame):
            self._sdf = data._sdf
        else:
            import pandas as pd
            pdf = data if isinstance(data, pd.DataFrame) else pd.DataFrame(data, columns=columns)
            self._sdf = _default_session().createDataFrame(pdf)

    @property
    def columns(self) -> List[str]:
        return self._sdf.columns

    def __getitem__(self, key):
        if isinstance(key, str):
            return Series(self._sdf[key], key)
        if isinstance(key, Series):
            return DataFrame(self._sdf.filter(key._expr))
        return DataFrame(self._sdf.select(*key))

    def __setitem__(self, key: str, value) -> None:
        expr = value._expr if isinstance(value, Series) else value
        if not isinstance(expr, Column):
            from raydp_b200.sql.functions import lit
            expr = lit(expr)
        self._sdf = self._sdf.withColumn(key, expr)

    def __getattr__(self, name: str):
        if name.startswith("_"):
            raise AttributeError(name)
        if name in self._sdf.columns:
            return self[name]
        raise AttributeError(name)

    def __len__(self) -> int:
        return self._sdf.count()

    @property
    def shape(self):
        return (len(self), len(self.columns))

    def to_spark(self, index_col=None) -> _SparkDF:
        return self._sdf

    def to_pandas(self):
        return self._sdf.toPandas()

    def head(self, n: int = 5):
        return self._sdf.limit(n).toPandas()

    def astype(self, dtype) -> "DataFrame":
        out = self._sdf
        for c in self.columns:
            out = out.withColumn(c, Series(out[c]).astype(dtype)._expr)
        return DataFrame(out)


def _default_session():
    from raydp_b200.sql.session import SparkSession
    active = SparkSession.getActiveSession()
    return active if active is not None else SparkSession.builder.master("local[2]").getOrCreate()


def range(start: int, end: Optional[int] = None, step: int = 1, num_partitions: Optional[int] = None) -> DataFrame:  # noqa: A001
    return DataFrame(_default_session().range(start, end, step, num_partitions))


def from_pandas(pdf) -> DataFrame:
    return DataFrame(pdf)
