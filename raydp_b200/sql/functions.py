"""``pyspark.sql.functions`` subset used by the reference's examples and tests.

Parity: ``col, lit, abs, log, rand, hour, dayofmonth, dayofweek, weekofyear, month, quarter,
year, udf`` (reference ``examples/data_process.py:4,28-84``, ``test_tf.py:42-44``),
``count`` / ``desc`` / ``row_number`` / ``monotonically_increasing_id`` / ``posexplode(array())``
(DLRM notebook cell 6).
"""
from __future__ import annotations

import builtins
from typing import Callable, Optional, Sequence, Union

import pyarrow as pa

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
def to_timestamp(c, format=None):  # noqa: A002
    if format is not None:
        return Column("call", ("to_timestamp_fmt", (as_column(c), format)))
    return _fn("to_timestamp", c)
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
def _plain(v):
    """Parameters that must be constants: unwrap ``lit(...)``."""
    if isinstance(v, Column):
        if v._kind != "lit":
            raise TypeError("this parameter must be a literal")
        return v._args[0]
    return v


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
def countDistinct(c, *cols): return count_distinct(c, *cols)
def first(c, ignorenulls: bool = False):
    # ignorenulls=False (the default) returns the first ROW's value even when it is NULL, as Spark does
    col = as_column(c)
    return Column("agg", ("first_value", col, bool(_plain(ignorenulls))), f"first({col._output_name()})")


def row_number() -> Column:
    return Column("rank_fn", ("row_number",))


def posexplode(c) -> Column:
    return Column("posexplode", (as_column(c),))


def explode(c) -> Column:
    return Column("explode", (as_column(c),))


# ---------------------------------------------------------------------------------------------------
# the wider pyspark.sql.functions surface (implementations: sql/fnlib.py, sql/aggregates.py,
# sql/windowing.py)
# ---------------------------------------------------------------------------------------------------
def _lit_or_col(v) -> Column:
    return v if isinstance(v, Column) else Column("lit", (v,))


def _unary(name: str):
    def fn(col):
        return _fn(name, col)
    fn.__name__ = name
    return fn


for _n in ("acos", "asin", "atan", "cosh", "sinh", "tan", "tanh", "acosh", "asinh", "atanh", "cbrt", "degrees",
           "radians", "expm1", "rint", "signum", "sign", "factorial", "bin", "hex", "unhex", "isnan", "initcap",
           "soundex", "ascii", "char_length", "character_length", "octet_length", "bit_length", "base64", "unbase64",
           "md5", "sha1", "crc32", "reverse", "last_day", "array_distinct", "array_max", "array_min", "array_sort",
           "flatten", "size", "cardinality", "array_size", "array_compact", "shuffle", "map_keys", "map_values",
           "map_entries", "map_from_entries", "sentences", "typeof", "bitwise_not"):
    globals()[_n] = _unary(_n)
del _n
bitwiseNOT = bitwise_not  # noqa: F821
toDegrees, toRadians = degrees, radians  # noqa: F821


def column(name: str) -> Column:
    return col(name)


def broadcast(df):
    """Join hint; the engine already broadcasts the smaller side, so the frame is returned as is."""
    return df


def atan2(a, b): return _fn("atan2", _num_or_col(a), _num_or_col(b))
def hypot(a, b): return _fn("hypot", _num_or_col(a), _num_or_col(b))
def pmod(a, b): return _fn("pmod", _num_or_col(a), _num_or_col(b))
def bround(c, scale: int = 0): return Column("call", ("bround", (as_column(c), scale)))
def conv(c, fromBase: int, toBase: int): return Column("call", ("conv", (as_column(c), fromBase, toBase)))
def shiftleft(c, numBits: int): return Column("call", ("shiftleft", (as_column(c), numBits)))
def shiftright(c, numBits: int): return Column("call", ("shiftright", (as_column(c), numBits)))


def _num_or_col(v) -> Column:
    return as_column(v) if isinstance(v, (str, Column)) else Column("lit", (v,))


def nanvl(a, b): return _fn("nanvl", a, b)
def nvl(a, b): return _fn("coalesce", a, b)
ifnull = nvl
def nvl2(a, b, c): return _fn("nvl2", a, b, c)
def nullif(a, b): return _fn("nullif", a, b)
def isnotnull(c): return as_column(c).isNotNull()
def equal_null(a, b): return _fn("eq_null_safe", a, b)


# strings ---------------------------------------------------------------------------------------------
def concat_ws(sep: str, *cols):
    cols = cols[0] if len(cols) == 1 and isinstance(cols[0], (list, tuple)) else cols
    return Column("call", ("concat_ws", (sep,) + tuple(as_column(c) for c in cols)))


def substring(c, pos: int, len: int):  # noqa: A002
    return Column("call", ("substring", (as_column(c), pos, len)))


substr = substring
def left(c, n): return Column("call", ("left", (as_column(c), _plain(n))))
def right(c, n): return Column("call", ("right", (as_column(c), _plain(n))))
def substring_index(c, delim: str, count: int): return Column("call", ("substring_index", (as_column(c), delim, count)))


def _plain(v):
    """Parameters that must be constants: unwrap ``lit(...)``."""
    if isinstance(v, Column):
        if v._kind != "lit":
            raise TypeError("this parameter must be a literal")
        return v._args[0]
    return v


def trim(c, trim=None): return Column("call", ("trim", (as_column(c), _plain(trim))))  # noqa: A002
def ltrim(c, trim=None): return Column("call", ("ltrim", (as_column(c), _plain(trim))))  # noqa: A002
def rtrim(c, trim=None): return Column("call", ("rtrim", (as_column(c), _plain(trim))))  # noqa: A002
def lpad(c, len: int, pad: str = " "): return Column("call", ("lpad", (as_column(c), len, pad)))  # noqa: A002
def rpad(c, len: int, pad: str = " "): return Column("call", ("rpad", (as_column(c), len, pad)))  # noqa: A002
def repeat(c, n: int): return Column("call", ("repeat", (as_column(c), _plain(n))))
def instr(c, substr: str): return Column("call", ("instr", (as_column(c), substr)))  # noqa: A002
def locate(substr: str, c, pos: int = 1): return Column("call", ("locate", (substr, as_column(c), pos)))  # noqa: A002
def translate(c, matching: str, replace: str): return Column("call", ("translate", (as_column(c), matching, replace)))
def regexp_extract(c, pattern: str, idx: int = 1): return Column("call", ("regexp_extract", (as_column(c), _plain(pattern), idx)))
def regexp_substr(c, pattern): return Column("call", ("regexp_substr", (as_column(c), _plain(pattern))))
def regexp_count(c, pattern): return Column("call", ("regexp_count", (as_column(c), _plain(pattern))))


def regexp_replace(c, pattern, replacement):
    return Column("call", ("regexp_replace", (as_column(c), _plain(pattern), _plain(replacement))))


def rlike(c, pattern): return Column("call", ("rlike", (as_column(c), _plain(pattern))))
regexp = regexp_like = rlike
def like(c, pattern): return Column("call", ("like", (as_column(c), _plain(pattern))))
def ilike(c, pattern): return Column("call", ("ilike", (as_column(c), _plain(pattern))))
def startswith(c, prefix): return as_column(c).startswith(prefix)
def endswith(c, suffix): return as_column(c).endswith(suffix)
def contains(c, other): return as_column(c).contains(other)


def split(c, pattern: str, limit: int = -1):
    return Column("call", ("split", (as_column(c), _plain(pattern), limit)))


def split_part(c, delimiter, partNum): return Column("call", ("split_part", (as_column(c), _plain(delimiter), _plain(partNum))))
def format_number(c, d: int): return Column("call", ("format_number", (as_column(c), d)))
def format_string(format: str, *cols): return Column("call", ("format_string", (format,) + tuple(as_column(c) for c in cols)))  # noqa: A002
printf = format_string
def levenshtein(left, right): return _fn("levenshtein", left, right)  # noqa: A002
def encode(c, charset: str): return Column("call", ("encode", (as_column(c), charset)))
def decode(c, charset: str): return Column("call", ("decode", (as_column(c), charset)))
def sha2(c, numBits: int): return Column("call", ("sha2", (as_column(c), numBits)))
sha = sha1  # noqa: F821
def xxhash64(*cols): return _fn("xxhash64", *cols)


def overlay(src, replace, pos, len=-1):  # noqa: A002
    return Column("call", ("overlay", (as_column(src), as_column(replace), _plain(pos), _plain(len))))


# dates -----------------------------------------------------------------------------------------------
def current_date(): return Column("call", ("current_date", ()))
curdate = current_date
def current_timestamp(): return Column("call", ("current_timestamp", ()))
now = localtimestamp = current_timestamp


def to_date(c, format: Optional[str] = None):  # noqa: A002
    return Column("call", ("to_date", (as_column(c), format)))


def date_add(start, days): return Column("call", ("date_add", (as_column(start), _lit_or_col(days) if not isinstance(days, str) else as_column(days))))
def date_sub(start, days): return Column("call", ("date_sub", (as_column(start), _lit_or_col(days) if not isinstance(days, str) else as_column(days))))
dateadd = date_add
def datediff(end, start): return _fn("datediff", end, start)
date_diff = datediff
def add_months(start, months): return Column("call", ("add_months", (as_column(start), _lit_or_col(months) if not isinstance(months, str) else as_column(months))))
def months_between(date1, date2, roundOff: bool = True): return Column("call", ("months_between", (as_column(date1), as_column(date2), roundOff)))
def next_day(date, dayOfWeek: str): return Column("call", ("next_day", (as_column(date), dayOfWeek)))
def date_format(date, format: str): return Column("call", ("date_format", (as_column(date), format)))  # noqa: A002
def date_trunc(format: str, timestamp): return Column("call", ("date_trunc", (format, as_column(timestamp))))  # noqa: A002
def trunc(date, format: str): return Column("call", ("trunc", (as_column(date), format)))  # noqa: A002
def from_unixtime(timestamp, format: str = "yyyy-MM-dd HH:mm:ss"): return Column("call", ("from_unixtime", (as_column(timestamp), format)))  # noqa: A002
def timestamp_seconds(c): return _fn("timestamp_seconds", c)
def make_date(year, month, day): return _fn("make_date", year, month, day)
def to_utc_timestamp(timestamp, tz: str): return Column("call", ("to_utc_timestamp", (as_column(timestamp), _plain(tz))))
def from_utc_timestamp(timestamp, tz: str): return Column("call", ("from_utc_timestamp", (as_column(timestamp), _plain(tz))))


def _duration_micros(text: str) -> int:
    units = {"microsecond": 1, "millisecond": 1000, "second": 10 ** 6, "minute": 60 * 10 ** 6, "hour": 3600 * 10 ** 6,
             "day": 86400 * 10 ** 6, "week": 7 * 86400 * 10 ** 6}
    total, parts = 0, text.strip().split()
    for qty, unit in zip(parts[0::2], parts[1::2]):
        total += int(qty) * units[unit.lower().rstrip("s")]
    return total


def window(timeColumn, windowDuration: str, slideDuration: Optional[str] = None, startTime: Optional[str] = None):
    """Time windows: a struct ``{start, end}`` per row for tumbling windows; with a ``slideDuration`` every
    window containing the timestamp (the row is repeated once per window, like Spark's Expand)."""
    width = _duration_micros(_plain(windowDuration))
    offset = _duration_micros(_plain(startTime)) if startTime else 0
    slide = _duration_micros(_plain(slideDuration)) if slideDuration is not None else width
    if slide > width:
        raise ValueError("The slide duration must be less than or equal to the window duration")
    # rows without a timestamp belong to no window and drop out, as in Spark
    windows = Column("call", ("sliding_windows", (as_column(timeColumn), width, slide, offset)))
    return Column("explode", (windows,), "window")


# arrays / maps / structs / json ---------------------------------------------------------------------------
def array_contains(c, value): return Column("call", ("array_contains", (as_column(c), _lit_or_col(value))))
def array_join(c, delimiter: str, null_replacement: Optional[str] = None):
    args = (as_column(c), delimiter) + ((null_replacement,) if null_replacement is not None else ())
    return Column("call", ("array_join", args))
def sort_array(c, asc: bool = True): return Column("call", ("sort_array", (as_column(c), asc)))  # noqa: A002
def element_at(c, extraction): return Column("call", ("element_at", (as_column(c), _lit_or_col(extraction))))
try_element_at = element_at
def get(c, index): return Column("call", ("get_item", (as_column(c), _lit_or_col(index))))
def slice(c, start, length): return Column("call", ("slice", (as_column(c), _lit_or_col(start), _lit_or_col(length))))  # noqa: A001
def array_position(c, value): return Column("call", ("array_position", (as_column(c), _lit_or_col(value))))
def array_remove(c, element): return Column("call", ("array_remove", (as_column(c), _lit_or_col(element))))
def array_union(a, b): return _fn("array_union", a, b)
def array_intersect(a, b): return _fn("array_intersect", a, b)
def array_except(a, b): return _fn("array_except", a, b)
def arrays_overlap(a, b): return _fn("arrays_overlap", a, b)
def array_append(c, value): return Column("call", ("array_append", (as_column(c), _lit_or_col(value))))
def array_prepend(c, value): return Column("call", ("array_prepend", (as_column(c), _lit_or_col(value))))
def array_insert(c, pos, value): return Column("call", ("array_insert", (as_column(c), _lit_or_col(pos), _lit_or_col(value))))
def array_repeat(c, count): return Column("call", ("array_repeat", (as_column(c), _lit_or_col(count))))
def arrays_zip(*cols): return _fn("arrays_zip", *cols)


def sequence(start, stop, step=None):
    args = (as_column(start), as_column(stop)) + ((as_column(step),) if step is not None else ())
    return Column("call", ("sequence", args))


def struct(*cols):
    cols = cols[0] if len(cols) == 1 and isinstance(cols[0], (list, tuple)) else cols
    exprs = [as_column(c) for c in cols]
    return Column("call", ("struct", (tuple(e._output_name() for e in exprs),) + tuple(exprs)))


def named_struct(*cols):
    names = [_plain(c) for c in cols[0::2]]
    return Column("call", ("struct", (tuple(names),) + tuple(as_column(c) for c in cols[1::2])))


def create_map(*cols):
    cols = cols[0] if len(cols) == 1 and isinstance(cols[0], (list, tuple)) else cols
    return _fn("create_map", *cols)


def map_from_arrays(keys, values): return _fn("map_from_arrays", keys, values)
def map_concat(*cols): return _fn("map_concat", *cols)
def map_contains_key(c, value): return Column("call", ("map_contains_key", (as_column(c), _lit_or_col(value))))
def get_json_object(c, path: str): return Column("call", ("get_json_object", (as_column(c), path)))


def json_tuple(c, *fields):
    return [Column("call", ("json_tuple_field", (as_column(c), f)), f"c{i}") for i, f in enumerate(fields)]


def from_json(c, schema, options=None):
    schema = _plain(schema)
    if isinstance(schema, str):
        try:
            schema = T.parse_type(schema)
        except ValueError:
            schema = T.parse_ddl(schema)
    arrow = schema.arrow if not isinstance(schema, T.StructType) else pa.struct(
        [(f.name, f.dataType.arrow) for f in schema.fields])
    return Column("call", ("from_json", (as_column(c), arrow)))


def to_json(c, options=None): return _fn("to_json", c)
def schema_of_json(json, options=None): return Column("call", ("schema_of_json", (_plain(json),)))  # noqa: A002


def explode_outer(c): return Column("explode", (as_column(c), True))
def posexplode_outer(c): return Column("posexplode", (as_column(c), True))


# higher-order functions: the lambda is applied once to placeholder columns -------------------------------
_LAMBDA_IDS = iter(range(1, 1 << 62))


def _lambda(f: Callable, arity_names: Sequence[str]):
    import inspect
    nargs = len(inspect.signature(f).parameters)
    uid = next(_LAMBDA_IDS)
    names = [f"__lambda_{nm}_{uid}" for nm in arity_names[:nargs]]
    body = f(*[Column("col", (nm,)) for nm in names])
    return _lit_or_col(body), tuple(names)


def transform(c, f: Callable):
    body, names = _lambda(f, ("x", "i"))
    return Column("hof", ("transform", (as_column(c),), body, names, ()))


def filter(c, f: Callable):  # noqa: A001
    body, names = _lambda(f, ("x", "i"))
    return Column("hof", ("filter", (as_column(c),), body, names, ()))


def exists(c, f: Callable):
    body, names = _lambda(f, ("x",))
    return Column("hof", ("exists", (as_column(c),), body, names, ()))


def forall(c, f: Callable):
    body, names = _lambda(f, ("x",))
    return Column("hof", ("forall", (as_column(c),), body, names, ()))


def zip_with(left, right, f: Callable):  # noqa: A002
    body, names = _lambda(f, ("x", "y"))
    return Column("hof", ("zip_with", (as_column(left), as_column(right)), body, names, ()))


def aggregate(c, initialValue, merge: Callable, finish: Optional[Callable] = None):
    body, names = _lambda(merge, ("acc", "x"))
    extra = [("initial", _lit_or_col(initialValue))]
    if finish is not None:
        fin_body, fin_names = _lambda(finish, ("acc",))
        extra += [("finish", fin_body), ("finish_var", fin_names[0])]
    return Column("hof", ("aggregate", (as_column(c),), body, names, tuple(extra)))


reduce = aggregate


def map_filter(c, f: Callable):
    body, names = _lambda(f, ("k", "v"))
    return Column("hof", ("map_filter", (as_column(c),), body, names, ()))


def transform_values(c, f: Callable):
    body, names = _lambda(f, ("k", "v"))
    return Column("hof", ("transform_values", (as_column(c),), body, names, ()))


def transform_keys(c, f: Callable):
    body, names = _lambda(f, ("k", "v"))
    return Column("hof", ("transform_keys", (as_column(c),), body, names, ()))


# aggregates ------------------------------------------------------------------------------------------
def _agg_p(fn: str, c, *params) -> Column:
    return Column("agg", (fn, as_column(c)) + tuple(params))


def _agg2(fn: str, a, b) -> Column:
    packed = Column("call", ("struct", (("0", "1"), as_column(a), as_column(b))))
    return Column("agg", (fn, packed), f"{fn}({as_column(a)._output_name()}, {as_column(b)._output_name()})")


def collect_list(c): return _agg_p("collect_list", c)
array_agg = collect_list
def collect_set(c): return _agg_p("collect_set", c)
def last(c, ignorenulls: bool = False): return _agg_p("last", c, bool(_plain(ignorenulls)))
def any_value(c, ignoreNulls: bool = False): return _agg_p("any_value", c, bool(_plain(ignoreNulls)))
def first_value(c, ignoreNulls: bool = False): return _agg_p("first_value", c, bool(_plain(ignoreNulls)))
def last_value(c, ignoreNulls: bool = False): return _agg_p("last_value", c, bool(_plain(ignoreNulls)))
def nth_value(c, offset: int, ignoreNulls: bool = False): return _agg_p("nth_value", c, int(_plain(offset)))
def approx_count_distinct(c, rsd: Optional[float] = None): return _agg_p("approx_count_distinct", c)
approxCountDistinct = approx_count_distinct
def sum_distinct(c): return _agg_p("sum_distinct", c)
sumDistinct = sum_distinct
def stddev_samp(c): return _agg("stddev", c)
std = stddev_samp
def stddev_pop(c): return _agg_p("stddev_pop", c)
def var_samp(c): return _agg("variance", c)
def var_pop(c): return _agg_p("var_pop", c)
def skewness(c): return _agg_p("skewness", c)
def kurtosis(c): return _agg_p("kurtosis", c)
def product(c): return _agg_p("product", c)
def bool_and(c): return _agg_p("bool_and", c)
every = bool_and
def bool_or(c): return _agg_p("bool_or", c)
some = bool_or
def bit_and(c): return _agg_p("bit_and", c)
def bit_or(c): return _agg_p("bit_or", c)
def bit_xor(c): return _agg_p("bit_xor", c)
def count_if(c): return _agg_p("count_if", c)
def median(c): return _agg_p("median", c)
def mode(c): return _agg_p("mode", c)
def percentile(c, percentage, frequency=1): return _agg_p("percentile", c, _plain_seq(percentage))
def percentile_approx(c, percentage, accuracy: int = 10000): return _agg_p("percentile_approx", c, _plain_seq(percentage))
approx_percentile = percentile_approx
def corr(a, b): return _agg2("corr", a, b)
def covar_pop(a, b): return _agg2("covar_pop", a, b)
def covar_samp(a, b): return _agg2("covar_samp", a, b)
def min_by(c, ord): return _agg2("min_by", c, ord)  # noqa: A002
def max_by(c, ord): return _agg2("max_by", c, ord)  # noqa: A002


def _plain_seq(v):
    if isinstance(v, (list, tuple)):
        return tuple(_plain(x) for x in v)
    return _plain(v)


def count_distinct(c, *cols):
    if cols:
        return _agg("count_distinct", struct(c, *cols))
    return _agg("count_distinct", c)


# second batch of scalar names ----------------------------------------------------------------------------
for _n in ("cot", "csc", "sec", "bit_count", "char", "json_array_length", "json_object_keys", "to_csv", "unix_date",
           "unix_micros", "unix_millis", "unix_seconds", "timestamp_millis", "timestamp_micros", "date_from_unix_date",
           "weekday", "to_number"):
    globals()[_n] = _unary(_n)
del _n
ceiling, ln, power, lcase, ucase, negative, negate, day = ceil, log, pow, lower, upper, (lambda c: -as_column(c)), \
    (lambda c: -as_column(c)), dayofmonth
positive = lambda c: as_column(c)  # noqa: E731
btrim = trim
try_to_number = to_number  # noqa: F821
to_unix_timestamp = unix_timestamp


def pi(): return lit(3.141592653589793)
def e(): return lit(2.718281828459045)
def uuid(): return Column("call", ("uuid", ()))
def current_database(): return lit("default")
current_schema = current_database
def current_catalog(): return lit("spark_catalog")
def current_user(): return lit(__import__("getpass").getuser())
def current_timezone(): return lit("UTC")
def input_file_name(): return lit("")
def getbit(c, pos): return _fn("getbit", c, _lit_or_col(pos))
bit_get = getbit
def shiftrightunsigned(c, numBits: int): return Column("call", ("shiftrightunsigned", (as_column(c), _plain(numBits))))
def elt(*cols): return _fn("elt", *[_lit_or_col(c) if not isinstance(c, str) else as_column(c) for c in cols])
def find_in_set(s, str_array): return _fn("find_in_set", s, str_array)
def position(substr, s, start=None): return Column("call", ("position", (_lit_or_col(substr), as_column(s)) + ((_plain(start),) if start is not None else ())))  # noqa: A002
def replace(src, search, replace=None): return Column("call", ("replace", (as_column(src), _plain(search)) + ((_plain(replace),) if replace is not None else ())))
def mask(c, upperChar="X", lowerChar="x", digitChar="n", otherChar=None): return Column("call", ("mask", (as_column(c), _plain(upperChar), _plain(lowerChar), _plain(digitChar), _plain(otherChar))))
def regexp_extract_all(c, regexp, idx=1): return Column("call", ("regexp_extract_all", (as_column(c), _plain(regexp), _plain(idx))))
def regexp_instr(c, regexp, idx=None): return Column("call", ("regexp_instr", (as_column(c), _plain(regexp))))
def str_to_map(text, pairDelim=",", keyValueDelim=":"): return Column("call", ("str_to_map", (as_column(text), _plain(pairDelim), _plain(keyValueDelim))))
def raise_error(errMsg): return Column("call", ("raise_error", (_lit_or_col(errMsg),)))
def width_bucket(v, min, max, numBucket): return _fn("width_bucket", v, _lit_or_col(min), _lit_or_col(max), _lit_or_col(numBucket))  # noqa: A002
def make_timestamp(years, months, days, hours, mins, secs, timezone=None): return _fn("make_timestamp", years, months, days, hours, mins, secs)
def date_part(field, source): return Column("call", ("date_part", (_plain(field), as_column(source))))
datepart = extract = date_part
def convert_timezone(sourceTz, targetTz, sourceTs): return Column("call", ("convert_timezone", (_plain(sourceTz), _plain(targetTz), as_column(sourceTs))))
def to_char(c, format): return date_format(c, _plain(format))  # noqa: A002
def try_add(a, b): return _fn("try_add", a, b)
def try_subtract(a, b): return _fn("try_subtract", a, b)
def try_multiply(a, b): return _fn("try_multiply", a, b)
def try_divide(a, b): return _fn("try_divide", a, b)
def try_sum(c): return sum(c)
def call_udf(udfName: str, *cols):
    from raydp_b200.sql.session import SparkSession
    return SparkSession.active()._registered_udfs[udfName](*cols)


def stack(n, *cols):
    """``stack(n, v1, ..., vk)``: ``n`` rows of ``k / n`` columns each (a generator, like ``explode``)."""
    rows = int(_plain(n))
    values = [_lit_or_col(c) if not isinstance(c, str) else as_column(c) for c in cols]
    width = -(-len(values) // rows)
    structs = []
    for r in range(rows):
        chunk = values[r * width:(r + 1) * width]
        chunk = chunk + [lit(None)] * (width - len(chunk))
        structs.append(struct(*[c.alias(f"col{i}") for i, c in enumerate(chunk)]))
    return Column("inline", (array(*structs),))


def inline(c): return Column("inline", (as_column(c),))
def inline_outer(c): return Column("inline", (as_column(c), True))


# window functions -----------------------------------------------------------------------------------------
def rank() -> Column: return Column("rank_fn", ("rank",))
def dense_rank() -> Column: return Column("rank_fn", ("dense_rank",))
def percent_rank() -> Column: return Column("rank_fn", ("percent_rank",))
def cume_dist() -> Column: return Column("rank_fn", ("cume_dist",))
def ntile(n: int) -> Column: return Column("rank_fn", ("ntile", int(_plain(n))))
def lag(c, offset: int = 1, default=None) -> Column: return Column("rank_fn", ("lag", as_column(c), int(_plain(offset)), default))
def lead(c, offset: int = 1, default=None) -> Column: return Column("rank_fn", ("lead", as_column(c), int(_plain(offset)), default))


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


class PandasUDFType:
    SCALAR, SCALAR_ITER, GROUPED_MAP, GROUPED_AGG = 200, 204, 201, 202


def pandas_udf(f=None, returnType=None, functionType=None):
    """``@pandas_udf("double")``, ``@pandas_udf(returnType=...)`` and ``pandas_udf(fn, "double")``: the
    function gets one ``pandas.Series`` per argument column (a whole partition at a time) and returns a
    Series / array of the same length."""
    if callable(f) and not isinstance(f, T.DataType):
        return UserDefinedFunction(_series_adapter(f), returnType if returnType is not None else T.DoubleType(),
                                   vectorized=True)
    rtype = f if f is not None else returnType
    return lambda fn: UserDefinedFunction(_series_adapter(fn), rtype if rtype is not None else T.DoubleType(),
                                          vectorized=True)


def _series_adapter(fn: Callable) -> Callable:
    def call(*arrays):
        import numpy as np
        import pandas as pd
        out = fn(*[pd.Series(a) for a in arrays])
        if isinstance(out, pd.Series):
            if out.dtype == object or str(out.dtype) in ("string", "str"):
                return np.array([None if v is None or v is pd.NA or (isinstance(v, float) and v != v) else v
                                 for v in out], dtype=object)
            return out.to_numpy()
        return np.asarray(out)
    call.__name__ = getattr(fn, "__name__", "pandas_udf")
    return call


def expr(text: str) -> Column:
    """SQL expression string -> Column (``F.expr("a + b * 2 AS c")``)."""
    from raydp_b200.sql.sqlparser import parse_expression
    return parse_expression(text)
