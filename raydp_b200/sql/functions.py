"""``pyspark.sql.functions`` subset used by the reference's examples and tests.

Parity: ``col, lit, abs, log, rand, hour, dayofmonth, dayofweek, weekofyear, month, quarter,
year, udf`` (reference ``examples/data_process.py:4,28-84``, ``test_tf.py:42-44``),
``count`` / ``desc`` / ``row_number`` / ``monotonically_increasing_id`` / ``posexplode(array())``
(DLRM notebook cell 6).
"""
from __future__ import annotations

import builtins
from typing import Callable, Optional, Union

from raydp_b200.sql import types as T
from raydp_b200.sql.column import Column, as_column


def col(name: str) -> Column:
    return Column("col", (name,))


column = col


def lit(value) -> Column:
    return value if isinstance(value, Column) else Column("lit", (value,))


def _fn(name: str, *args) -> Column:
    return Column("call", (name, tuple(as_column(a) if isinstance(a, (Column, str)) else a for a in args)))


def abs(c) -> Column:  # noqa: A001 - mirrors pyspark
    return _fn("abs", c)


def log(arg1, arg2=None) -> Column:
    if arg2 is None:
        return _fn("log", arg1)
    return Column("call", ("log", (lit(float(arg1)), as_column(arg2))))


def log1p(c): return _fn("log1p", c)
def log2(c): return _fn("log2", c)
def log10(c): return _fn("log10", c)
def exp(c): return _fn("exp", c)
def sqrt(c): return _fn("sqrt", c)
def floor(c): return _fn("floor", c)
def ceil(c): return _fn("ceil", c)
def sin(c): return _fn("sin", c)
def cos(c): return _fn("cos", c)


def pow(c, p) -> Column:  # noqa: A001
    return Column("call", ("pow", (as_column(c), p)))


def round(c, scale: int = 0) -> Column:  # noqa: A001
    return Column("call", ("round", (as_column(c), scale)))


def rand(seed: Optional[int] = None) -> Column:
    return Column("call", ("rand", (seed,)))


def randn(seed: Optional[int] = None) -> Column:
    return Column("call", ("randn", (seed,)))


def monotonically_increasing_id() -> Column:
    return Column("call", ("monotonically_increasing_id", ()))


def spark_partition_id() -> Column:
    return Column("call", ("spark_partition_id", ()))


def hour(c): return _fn("hour", c)
def minute(c): return _fn("minute", c)
def second(c): return _fn("second", c)
def dayofmonth(c): return _fn("dayofmonth", c)
def dayofweek(c): return _fn("dayofweek", c)
def dayofyear(c): return _fn("dayofyear", c)
def weekofyear(c): return _fn("weekofyear", c)
def month(c): return _fn("month", c)
def quarter(c): return _fn("quarter", c)
def year(c): return _fn("year", c)
def to_timestamp(c, fmt=None): return _fn("to_timestamp", c)
def unix_timestamp(c, fmt=None): return _fn("unix_timestamp", c)
def coalesce(*cols): return _fn("coalesce", *cols)
def concat(*cols): return _fn("concat", *cols)
def upper(c): return _fn("upper", c)
def lower(c): return _fn("lower", c)
def length(c): return _fn("length", c)
def greatest(*cols): return _fn("greatest", *cols)
def least(*cols): return _fn("least", *cols)
def isnull(c): return as_column(c).isNull()
def array(*cols): return _fn("array", *(cols[0] if len(cols) == 1 and isinstance(cols[0], (list, tuple)) else cols))


def hash(*cols) -> Column:  # noqa: A001
    return _fn("hash", *cols)


def when(condition: Column, value) -> Column:
    return Column("when", (((condition, lit(value)),), None))


def desc(c) -> Column:
    return as_column(c).desc()


def asc(c) -> Column:
    return as_column(c).asc()


# aggregate expressions (used with DataFrame.agg / GroupedData.agg)
def _agg(fn: str, c) -> Column:
    target = None if (isinstance(c, str) and c == "*") else as_column(c)
    return Column("agg", (fn, target))


def count(c="*"): return _agg("count", c)
def sum(c): return _agg("sum", c)  # noqa: A001
def avg(c): return _agg("mean", c)
mean = avg
def min(c): return _agg("min", c)  # noqa: A001
def max(c): return _agg("max", c)  # noqa: A001
def stddev(c): return _agg("stddev", c)
def variance(c): return _agg("variance", c)
def countDistinct(c): return _agg("count_distinct", c)
def first(c): return _agg("first", c)


def row_number() -> Column:
    return Column("rank_fn", ("row_number",))


def posexplode(c) -> Column:
    return Column("posexplode", (as_column(c),))


def explode(c) -> Column:
    return Column("explode", (as_column(c),))


class UserDefinedFunction:
    def __init__(self, func: Callable, returnType=T.StringType(), vectorized: bool = False):
        self.func = func
        self.returnType = T.parse_type(returnType)
        self.vectorized = vectorized
        self.__name__ = getattr(func, "__name__", "udf")

    def __call__(self, *cols) -> Column:
        args = tuple(as_column(c) for c in cols)
        return Column("vudf" if self.vectorized else "udf", (self.func, self.returnType, args))


def udf(f: Union[Callable, str, T.DataType, None] = None, returnType=T.StringType()):
    """``udf(fn, "int")``, ``@udf("int")`` and ``@udf`` forms, like pyspark."""
    if f is None or isinstance(f, (str, T.DataType)):
        rtype = f if f is not None else returnType
        return lambda fn: UserDefinedFunction(fn, rtype)
    return UserDefinedFunction(f, returnType)


def vectorized_udf(returnType=T.DoubleType()):
    """Numpy-vectorised UDF: the function receives numpy arrays (built-in engine extension)."""
    return lambda fn: UserDefinedFunction(fn, returnType, vectorized=True)


pandas_udf = vectorized_udf


def expr(text: str) -> Column:
    """SQL expression string -> Column (``F.expr("a + b * 2 AS c")``)."""
    from raydp_b200.sql.sqlparser import parse_expression
    return parse_expression(text)
