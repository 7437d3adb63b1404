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


_DTYPES = {"float": "double", "float64": "double", "float32": "float", "int": "bigint", "int64": "bigint",
           "int32": "int", "int16": "smallint", "int8": "tinyint", "str": "string", "object": "string",
           "bool": "boolean", "datetime64[ns]": "timestamp", "datetime64": "timestamp"}


def _spark_type(dtype) -> str:
    name = getattr(dtype, "__name__", str(dtype)).lower()
    return _DTYPES.get(name, name)


class Series:
    """A column expression, bound to the frame it came from so that reductions can run."""

    def __init__(self, expr: Column, name: Optional[str] = None, frame: Optional["DataFrame"] = None):
        self._expr = expr
        self.name = name
        self._frame = frame

    def _wrap(self, expr: Column) -> "Series":
        return Series(expr, self.name, self._frame)

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
    def __mod__(self, o): return self._wrap(self._expr % self._unwrap(o))
    def __pow__(self, o): return self._wrap(self._expr ** self._unwrap(o))
    def __neg__(self): return self._wrap(-self._expr)
    def __lt__(self, o): return self._wrap(self._expr < self._unwrap(o))
    def __le__(self, o): return self._wrap(self._expr <= self._unwrap(o))
    def __gt__(self, o): return self._wrap(self._expr > self._unwrap(o))
    def __ge__(self, o): return self._wrap(self._expr >= self._unwrap(o))
    def __eq__(self, o): return self._wrap(self._expr == self._unwrap(o))  # type: ignore[override]
    def __ne__(self, o): return self._wrap(self._expr != self._unwrap(o))  # type: ignore[override]
    def __and__(self, o): return self._wrap(self._expr & self._unwrap(o))
    def __or__(self, o): return self._wrap(self._expr | self._unwrap(o))
    def __invert__(self): return self._wrap(~self._expr)
    __hash__ = object.__hash__

    def astype(self, dtype) -> "Series":
        return self._wrap(self._expr.cast(_spark_type(dtype)))

    def rename(self, name: str) -> "Series":
        return Series(self._expr, name, self._frame)

    def abs(self): return self._wrap(_F().abs(self._expr))
    def round(self, decimals: int = 0): return self._wrap(_F().round(self._expr, decimals))
    def isnull(self): return self._wrap(self._expr.isNull())
    isna = isnull
    def notnull(self): return self._wrap(self._expr.isNotNull())
    notna = notnull
    def isin(self, values): return self._wrap(self._expr.isin(list(values)))
    def between(self, left, right): return self._wrap(self._expr.between(left, right))

    def fillna(self, value):
        return self._wrap(_F().coalesce(self._expr, _F().lit(value)))

    def clip(self, lower=None, upper=None):
        expr = self._expr
        if lower is not None:
            expr = _F().greatest(expr, _F().lit(lower))
        if upper is not None:
            expr = _F().least(expr, _F().lit(upper))
        return self._wrap(expr)

    def apply(self, func, args=(), **kwds):
        """Element-wise Python function; the result type is taken from the function's return annotation
        (``-> float``), else string -- the same rule pandas-on-Spark applies."""
        import inspect
        hint = inspect.signature(func).return_annotation if callable(func) else inspect.Signature.empty
        rtype = _spark_type(hint) if hint is not inspect.Signature.empty else "string"
        return self._wrap(_F().udf(lambda v, _f=func: None if v is None else _f(v, *args, **kwds), rtype)(self._expr))

    def map(self, arg):  # noqa: A003
        if isinstance(arg, dict):
            out = None
            for k, v in arg.items():
                cond = self._expr == k
                out = _F().when(cond, _F().lit(v)) if out is None else out.when(cond, _F().lit(v))
            return self._wrap(out if out is not None else _F().lit(None))
        return self.apply(arg)

    @property
    def str(self):  # noqa: A003
        return _StrAccessor(self)

    @property
    def dt(self):
        return _DtAccessor(self)

    # ---- actions (need the frame)
    def _sdf(self):
        if self._frame is None:
            raise ValueError("this Series is not attached to a DataFrame")
        return self._frame._sdf.select(self._expr.alias(self.name or "0"))

    def _reduce(self, fn_name: str):
        agg = getattr(_F(), fn_name)(self._expr)
        return self._frame._sdf.agg(agg.alias("r")).collect()[0][0]

    def sum(self): return self._reduce("sum")  # noqa: A003
    def mean(self): return self._reduce("avg")
    def min(self): return self._reduce("min")  # noqa: A003
    def max(self): return self._reduce("max")  # noqa: A003
    def std(self): return self._reduce("stddev")
    def var(self): return self._reduce("variance")
    def median(self): return self._reduce("median")
    def count(self): return self._reduce("count")
    def nunique(self): return self._reduce("countDistinct")

    def unique(self):
        return self._sdf().distinct().toPandas().iloc[:, 0].to_numpy()

    def value_counts(self):
        name = self.name or "0"
        pdf = self._sdf().groupBy(name).count().toPandas()
        return pdf.set_index(name)["count"].sort_values(ascending=False)

    def to_pandas(self):
        return self._sdf().toPandas().iloc[:, 0]

    to_numpy = lambda self: self.to_pandas().to_numpy()  # noqa: E731
    tolist = to_list = lambda self: self.to_pandas().tolist()  # noqa: E731

    def head(self, n: int = 5):
        return self._sdf().limit(n).toPandas().iloc[:, 0]

    def __len__(self) -> int:
        return self._frame._sdf.count()

    def __repr__(self):
        return repr(self.head(10)) if self._frame is not None else f"Series<{self._expr!r}>"


def _F():
    from raydp_b200.sql import functions
    return functions


class _StrAccessor:
    def __init__(self, series: Series):
        self._s = series

    def upper(self): return self._s._wrap(_F().upper(self._s._expr))
    def lower(self): return self._s._wrap(_F().lower(self._s._expr))
    def len(self): return self._s._wrap(_F().length(self._s._expr))  # noqa: A003
    def strip(self): return self._s._wrap(_F().trim(self._s._expr))
    def contains(self, pat: str, regex: bool = True):
        return self._s._wrap(self._s._expr.rlike(pat) if regex else self._s._expr.contains(pat))
    def startswith(self, pat: str): return self._s._wrap(self._s._expr.startswith(pat))
    def endswith(self, pat: str): return self._s._wrap(self._s._expr.endswith(pat))
    def replace(self, pat: str, repl: str, regex: bool = True):
        import re
        return self._s._wrap(_F().regexp_replace(self._s._expr, pat if regex else re.escape(pat), repl))
    def slice(self, start=None, stop=None):  # noqa: A003
        start = start or 0
        return self._s._wrap(_F().substring(self._s._expr, start + 1, (stop - start) if stop is not None else 1 << 30))
    def split(self, pat: str = " "):
        import re
        return self._s._wrap(_F().split(self._s._expr, re.escape(pat)))


class _DtAccessor:
    def __init__(self, series: Series):
        self._s = series

    def _f(self, name): return self._s._wrap(getattr(_F(), name)(self._s._expr))
    year = property(lambda self: self._f("year"))
    month = property(lambda self: self._f("month"))
    day = property(lambda self: self._f("dayofmonth"))
    hour = property(lambda self: self._f("hour"))
    minute = property(lambda self: self._f("minute"))
    second = property(lambda self: self._f("second"))
    dayofyear = property(lambda self: self._f("dayofyear"))
    quarter = property(lambda self: self._f("quarter"))
    # pandas: Monday = 0; Spark's dayofweek: Sunday = 1
    dayofweek = property(lambda self: self._s._wrap((_F().dayofweek(self._s._expr) + 5) % 7))
    date = property(lambda self: self._s._wrap(_F().to_date(self._s._expr)))


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

    # ---- structure
    @property
    def columns(self) -> List[str]:
        return self._sdf.columns

    @columns.setter
    def columns(self, names) -> None:
        self._sdf = self._sdf.toDF(*list(names))

    @property
    def dtypes(self):
        import pandas as pd
        return pd.Series({n: t for n, t in self._sdf.dtypes})

    @property
    def shape(self):
        return (len(self), len(self.columns))

    @property
    def empty(self) -> bool:
        return self._sdf.isEmpty()

    @property
    def spark(self):
        return _SparkAccessor(self)

    @property
    def loc(self):
        return _Loc(self)

    def __len__(self) -> int:
        return self._sdf.count()

    def __repr__(self):
        return repr(self.head(10))

    def __getitem__(self, key):
        if isinstance(key, str):
            return Series(self._sdf[key], key, self)
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

    def copy(self, deep: bool = True) -> "DataFrame":
        return DataFrame(self._sdf)

    def reset_index(self, drop: bool = True, **_kw) -> "DataFrame":
        return DataFrame(self._sdf)

    # ---- conversions
    def to_spark(self, index_col=None) -> _SparkDF:
        return self._sdf

    def to_pandas(self):
        return self._sdf.toPandas()

    _to_pandas = to_pandas

    def to_numpy(self):
        return self.to_pandas().to_numpy()

    def to_parquet(self, path: str, mode: str = "overwrite", partition_cols=None, **_kw) -> None:
        writer = self._sdf.write.mode(mode)
        if partition_cols:
            writer = writer.partitionBy(*([partition_cols] if isinstance(partition_cols, str) else partition_cols))
        writer.parquet(path)

    def to_csv(self, path: Optional[str] = None, header: bool = True, mode: str = "overwrite", **_kw):
        if path is None:
            return self.to_pandas().to_csv(index=False, header=header)
        self._sdf.write.mode(mode).csv(path, header=header)

    def to_json(self, path: Optional[str] = None, mode: str = "overwrite", **_kw):
        if path is None:
            return self.to_pandas().to_json(orient="records")
        self._sdf.write.mode(mode).json(path)

    def head(self, n: int = 5):
        return self._sdf.limit(n).toPandas()

    def iterrows(self):
        return self.to_pandas().iterrows()

    # ---- column-wise transforms
    def astype(self, dtype) -> "DataFrame":
        out = self._sdf
        per_col = dtype if isinstance(dtype, dict) else {c: dtype for c in self.columns}
        for c, t in per_col.items():
            out = out.withColumn(c, out[c].cast(_spark_type(t)))
        return DataFrame(out)

    def assign(self, **kwargs) -> "DataFrame":
        out = self.copy()
        for k, v in kwargs.items():
            out[k] = v(out) if callable(v) else v
        return out

    def drop(self, labels=None, axis=1, columns=None, **_kw) -> "DataFrame":
        names = columns if columns is not None else labels
        names = [names] if isinstance(names, str) else list(names)
        return DataFrame(self._sdf.drop(*names))

    def rename(self, columns: Optional[dict] = None, **_kw) -> "DataFrame":
        return DataFrame(self._sdf.withColumnsRenamed(dict(columns or {})))

    def fillna(self, value, **_kw) -> "DataFrame":
        return DataFrame(self._sdf.fillna(value))

    def dropna(self, how: str = "any", subset=None, **_kw) -> "DataFrame":
        return DataFrame(self._sdf.dropna(how=how, subset=subset))

    def isnull(self) -> "DataFrame":
        return DataFrame(self._sdf.select(*[self._sdf[c].isNull().alias(c) for c in self.columns]))

    isna = isnull

    def drop_duplicates(self, subset=None, **_kw) -> "DataFrame":
        return DataFrame(self._sdf.dropDuplicates([subset] if isinstance(subset, str) else subset))

    def sort_values(self, by, ascending=True, **_kw) -> "DataFrame":
        by = [by] if isinstance(by, str) else list(by)
        return DataFrame(self._sdf.orderBy(*by, ascending=ascending))

    def sample(self, n=None, frac: Optional[float] = None, replace: bool = False, random_state=None) -> "DataFrame":
        if frac is None:
            raise ValueError("pandas-on-Spark needs frac= (a row count cannot be sampled exactly without a pass)")
        return DataFrame(self._sdf.sample(replace, frac, random_state))

    def query(self, expr: str) -> "DataFrame":
        return DataFrame(self._sdf.filter(expr.replace("==", "=")))

    def apply(self, func, axis=0, **_kw):
        """``axis=1``: ``func(row)`` with the row as a pandas Series; the result is a string column unless
        the function carries a return annotation (``-> float``)."""
        if axis in (1, "columns"):
            import inspect
            hint = inspect.signature(func).return_annotation
            rtype = _spark_type(hint) if hint is not inspect.Signature.empty else "string"
            names = self.columns

            def call(*values, _f=func, _n=tuple(names)):
                import pandas as pd
                return _f(pd.Series(list(values), index=list(_n)))
            return Series(_F().udf(call, rtype)(*names), None, self)
        return DataFrame(self._sdf.select(*[func(self[c])._expr.alias(c) for c in self.columns]))

    # ---- relational
    def merge(self, right: "DataFrame", how: str = "inner", on=None, left_on=None, right_on=None,
              suffixes=("_x", "_y")) -> "DataFrame":
        how = {"outer": "full"}.get(how, how)
        rs = right._sdf
        if on is None and left_on is None:
            on = [c for c in self.columns if c in right.columns]
        if on is not None:
            keys = [on] if isinstance(on, str) else list(on)
            left_sdf = self._sdf
            for c in rs.columns:
                if c in self.columns and c not in keys:
                    left_sdf = left_sdf.withColumnRenamed(c, c + suffixes[0])
                    rs = rs.withColumnRenamed(c, c + suffixes[1])
            return DataFrame(left_sdf.join(rs, keys, how))
        lk = [left_on] if isinstance(left_on, str) else list(left_on)
        rk = [right_on] if isinstance(right_on, str) else list(right_on)
        cond = None
        for a, b in zip(lk, rk):
            eq = _F().col(a) == _F().col(b)
            cond = eq if cond is None else cond & eq
        return DataFrame(self._sdf.join(rs, cond, how))

    def join(self, right: "DataFrame", on=None, how: str = "left", **_kw) -> "DataFrame":
        return self.merge(right, how=how, on=on)

    def groupby(self, by, as_index: bool = True, **_kw) -> "GroupBy":
        return GroupBy(self, [by] if isinstance(by, str) else list(by), as_index)

    # ---- reductions (numeric columns) -> pandas
    def _numeric(self) -> List[str]:
        return [n for n, t in self._sdf.dtypes if t in _NUMERIC]

    def _reduce(self, fn_name: str):
        import pandas as pd
        row = self._sdf.agg(*[getattr(_F(), fn_name)(c).alias(c) for c in self._numeric()]).collect()[0]
        return pd.Series(row.asDict())

    def sum(self): return self._reduce("sum")  # noqa: A003
    def mean(self): return self._reduce("avg")
    def min(self): return self._reduce("min")  # noqa: A003
    def max(self): return self._reduce("max")  # noqa: A003
    def std(self): return self._reduce("stddev")

    def count(self):
        import pandas as pd
        row = self._sdf.agg(*[_F().count(c).alias(c) for c in self.columns]).collect()[0]
        return pd.Series(row.asDict())

    def nunique(self):
        import pandas as pd
        row = self._sdf.agg(*[_F().countDistinct(c).alias(c) for c in self.columns]).collect()[0]
        return pd.Series(row.asDict())

    def describe(self):
        pdf = self._sdf.select(*self._numeric()).summary().toPandas().set_index("summary")
        return pdf.astype(float)


_NUMERIC = ("int", "bigint", "smallint", "tinyint", "float", "double")


class _SparkAccessor:
    def __init__(self, frame: DataFrame):
        self._frame = frame

    def frame(self, index_col=None):
        return self._frame._sdf

    def cache(self):
        self._frame._sdf.cache()
        return self._frame

    def repartition(self, n: int):
        return DataFrame(self._frame._sdf.repartition(n))


class _Loc:
    def __init__(self, frame: DataFrame):
        self._frame = frame

    def __getitem__(self, key):
        rows, cols = key if isinstance(key, tuple) else (key, None)
        out = self._frame
        if isinstance(rows, Series):
            out = out[rows]
        elif not (isinstance(rows, slice) and rows == slice(None)):
            raise NotImplementedError("loc supports boolean masks and ':' for rows (there is no index)")
        if cols is None or (isinstance(cols, slice) and cols == slice(None)):
            return out
        return out[cols] if isinstance(cols, str) else out[list(cols)]


class GroupBy:
    def __init__(self, frame: DataFrame, keys: List[str], as_index: bool = True, columns: Optional[List[str]] = None):
        self._frame, self._keys, self._as_index, self._columns = frame, keys, as_index, columns

    def __getitem__(self, cols):
        return GroupBy(self._frame, self._keys, self._as_index, [cols] if isinstance(cols, str) else list(cols))

    def _targets(self) -> List[str]:
        if self._columns is not None:
            return self._columns
        return [n for n, t in self._frame._sdf.dtypes if n not in self._keys and t in _NUMERIC]

    def _run(self, exprs) -> DataFrame:
        return DataFrame(self._frame._sdf.groupBy(*self._keys).agg(*exprs).orderBy(*self._keys))

    def _simple(self, spark_fn: str) -> DataFrame:
        return self._run([getattr(_F(), spark_fn)(c).alias(c) for c in self._targets()])

    def sum(self): return self._simple("sum")  # noqa: A003
    def mean(self): return self._simple("avg")
    def min(self): return self._simple("min")  # noqa: A003
    def max(self): return self._simple("max")  # noqa: A003
    def std(self): return self._simple("stddev")
    def first(self): return self._simple("first")
    def nunique(self): return self._simple("countDistinct")

    def count(self):
        cols = self._columns or [c for c in self._frame.columns if c not in self._keys]
        return self._run([_F().count(c).alias(c) for c in cols])

    def size(self):
        return self._run([_F().count("*").alias("size")])

    def agg(self, func=None, **named) -> DataFrame:
        names = {"mean": "avg", "std": "stddev", "var": "variance", "nunique": "countDistinct", "size": "count"}
        exprs = []
        if named:                                   # agg(total=("v", "sum"))
            for out, (col, fn) in named.items():
                exprs.append(getattr(_F(), names.get(fn, fn))(col).alias(out))
        elif isinstance(func, dict):
            for col, fns in func.items():
                for fn in ([fns] if isinstance(fns, str) else fns):
                    label = col if isinstance(fns, str) else f"{col}_{fn}"
                    exprs.append(getattr(_F(), names.get(fn, fn))(col).alias(label))
        else:
            for fn in ([func] if isinstance(func, str) else func):
                for col in self._targets():
                    label = col if isinstance(func, str) else f"{col}_{fn}"
                    exprs.append(getattr(_F(), names.get(fn, fn))(col).alias(label))
        return self._run(exprs)

    aggregate = agg

    def apply(self, func, *args, **kwargs) -> DataFrame:
        """``func(pandas.DataFrame) -> pandas.DataFrame`` per group."""
        import pandas as pd
        pdf = self._frame.to_pandas()
        pieces = [func(g.reset_index(drop=True), *args, **kwargs)
                  for _, g in pdf.groupby(self._keys, sort=True, dropna=False)]
        return DataFrame(pd.concat(pieces, ignore_index=True)) if pieces else self._frame


def _default_session():
    from raydp_b200.sql.session import SparkSession
    active = SparkSession.getActiveSession()
    return active if active is not None else SparkSession.builder.master("local[2]").getOrCreate()


def range(start: int, end: Optional[int] = None, step: int = 1, num_partitions: Optional[int] = None) -> DataFrame:  # noqa: A001
    return DataFrame(_default_session().range(start, end, step, num_partitions))


def from_pandas(pdf) -> DataFrame:
    return DataFrame(pdf)


def read_parquet(path, columns=None, **_kw) -> DataFrame:
    sdf = _default_session().read.parquet(path)
    return DataFrame(sdf.select(*columns) if columns else sdf)


def read_csv(path, sep: str = ",", header="infer", names=None, **_kw) -> DataFrame:
    reader = _default_session().read.option("header", header is not None).option("inferSchema", True).option("sep", sep)
    sdf = reader.csv(path)
    return DataFrame(sdf.toDF(*names) if names else sdf)


def read_json(path, **_kw) -> DataFrame:
    return DataFrame(_default_session().read.json(path))


def concat(objs, axis=0, ignore_index: bool = True, **_kw) -> DataFrame:
    frames = list(objs)
    out = frames[0]._sdf
    for f in frames[1:]:
        out = out.unionByName(f._sdf, allowMissingColumns=True)
    return DataFrame(out)


def sql(query: str, **_kw) -> DataFrame:
    return DataFrame(_default_session().sql(query))


def merge(left: DataFrame, right: DataFrame, **kw) -> DataFrame:
    return left.merge(right, **kw)


def to_datetime(arg, format=None, **_kw):  # noqa: A002
    if isinstance(arg, Series):
        from raydp_b200.sql import functions as F
        return arg._wrap(F.to_timestamp(arg._expr, _strftime_to_java(format) if format else None))
    import pandas as pd
    return pd.to_datetime(arg, format=format)


def _strftime_to_java(fmt: str) -> str:
    for c, java in (("%Y", "yyyy"), ("%m", "MM"), ("%d", "dd"), ("%H", "HH"), ("%M", "mm"), ("%S", "ss"), ("%y", "yy")):
        fmt = fmt.replace(c, java)
    return fmt


def get_dummies(data: DataFrame, columns=None, prefix_sep: str = "_", dtype=None, **_kw) -> DataFrame:
    """One-hot columns for the listed (or all string) columns; categories are found with one distinct pass."""
    from raydp_b200.sql import functions as F
    sdf = data._sdf
    cols = columns or [n for n, t in sdf.dtypes if t == "string"]
    out = sdf
    for c in cols:
        cats = sorted(r[0] for r in sdf.select(c).distinct().collect() if r[0] is not None)
        for cat in cats:
            out = out.withColumn(f"{c}{prefix_sep}{cat}", F.when(F.col(c) == cat, 1).otherwise(0).cast("int"))
        out = out.drop(c)
    return DataFrame(out)
