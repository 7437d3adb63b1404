"""Column expressions evaluated per partition with ``pyarrow.compute``.

Parity: the pyspark ``Column`` operators the reference's examples and tests use --
arithmetic / comparison on ``col(...)`` (reference ``examples/data_process.py:10-24,63-68``),
``alias``, ``cast``, ``isNull``, ``desc`` and python UDFs (``@udf("int")`` reference
``examples/data_process.py:35-49``).  An expression is a small picklable tree; executors
evaluate it against one Arrow table (a partition) at a time, so the hot loops are Arrow's C++
kernels, not Python.
"""
from __future__ import annotations

import datetime as _dt
from typing import Any, Callable, Dict, List, Optional, Sequence

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

from raydp_b200.sql import types as T


class EvalContext:
    """Per-partition evaluation state (partition index drives rand seeds and row ids)."""

    def __init__(self, partition_index: int = 0, session_tz: str = "UTC"):
        self.partition_index = partition_index
        self.session_tz = session_tz


def _to_array(value, n: int):
    if isinstance(value, (pa.Array, pa.ChunkedArray)):
        return value
    if isinstance(value, pa.Scalar):
        value = value.as_py()
    return pa.array([value] * n) if not isinstance(value, (int, float, bool)) or n < 0 else \
        pa.repeat(pa.scalar(value), n)


def _is_arrayish(v) -> bool:
    return isinstance(v, (pa.Array, pa.ChunkedArray))


class Column:
    """A node of the expression tree: ``kind`` + ``args`` (+ optional output name)."""

    def __init__(self, kind: str, args: tuple = (), name: Optional[str] = None):
        self._kind = kind
        self._args = args
        self._name = name

    # ------------------------------------------------------------------ naming
    def alias(self, *names: str, **_metadata) -> "Column":
        # generators (posexplode, map explode) take one alias per produced column
        return Column(self._kind, self._args, names[0] if len(names) == 1 else tuple(names))

    name = alias

    def _output_name(self) -> str:
        if self._name is not None:
            return self._name
        if self._kind == "col":
            return self._args[0]
        if self._kind == "lit":
            return str(self._args[0])
        if self._kind == "call":
            inner = ", ".join(a._output_name() if isinstance(a, Column) else str(a) for a in self._args[1])
            return f"{self._args[0]}({inner})"
        if self._kind == "binop":
            return f"({self._args[1]._output_name()} {self._args[0]} {self._args[2]._output_name()})"
        if self._kind in ("cast", "sort"):
            inner = self._args[0]._output_name()
            return inner if self._kind == "sort" or self._args[0]._kind == "col" else \
                f"CAST({inner} AS {self._args[1].simpleString().upper()})"
        if self._kind == "udf":
            return f"{self._args[0].__name__}(...)"
        if self._kind == "arrowfn":
            return f"{getattr(self._args[0], '__name__', 'fn')}(...)"
        return self._kind

    # ------------------------------------------------------------------ operators
    def _bin(self, op: str, other, reverse: bool = False) -> "Column":
        other = other if isinstance(other, Column) else Column("lit", (other,))
        return Column("binop", (op, other, self) if reverse else (op, self, other))

    def __add__(self, o): return self._bin("+", o)
    def __radd__(self, o): return self._bin("+", o, True)
    def __sub__(self, o): return self._bin("-", o)
    def __rsub__(self, o): return self._bin("-", o, True)
    def __mul__(self, o): return self._bin("*", o)
    def __rmul__(self, o): return self._bin("*", o, True)
    def __truediv__(self, o): return self._bin("/", o)
    def __rtruediv__(self, o): return self._bin("/", o, True)
    def __mod__(self, o): return self._bin("%", o)
    def __pow__(self, o): return self._bin("**", o)
    def __lt__(self, o): return self._bin("<", o)
    def __le__(self, o): return self._bin("<=", o)
    def __gt__(self, o): return self._bin(">", o)
    def __ge__(self, o): return self._bin(">=", o)
    def __eq__(self, o): return self._bin("==", o)  # type: ignore[override]
    def __ne__(self, o): return self._bin("!=", o)  # type: ignore[override]
    def __and__(self, o): return self._bin("&", o)
    def __or__(self, o): return self._bin("|", o)
    def __invert__(self): return Column("call", ("not", (self,)))
    def __neg__(self): return Column("call", ("negate", (self,)))
    __hash__ = object.__hash__

    def __bool__(self):
        raise ValueError("Cannot convert column into bool: use '&' for 'and', '|' for 'or', "
                         "'~' for 'not' when building DataFrame boolean expressions.")

    def cast(self, data_type) -> "Column":
        return Column("cast", (self, T.parse_type(data_type)), self._name)

    astype = cast

    def isNull(self) -> "Column":
        return Column("call", ("is_null", (self,)))

    def isNotNull(self) -> "Column":
        return Column("call", ("is_valid", (self,)))

    def isin(self, *values) -> "Column":
        if len(values) == 1 and isinstance(values[0], (list, tuple, set)):
            values = tuple(values[0])
        return Column("isin", (self, tuple(values)))

    def between(self, lo, hi) -> "Column":
        return (self >= lo) & (self <= hi)

    def desc(self) -> "Column":
        return Column("sort", (self, "descending"))

    def asc(self) -> "Column":
        return Column("sort", (self, "ascending"))

    # Arrow sorts nulls last in either direction; the placement is kept as a third field for the sorters
    def asc_nulls_first(self) -> "Column":
        return Column("sort", (self, "ascending", "at_start"))

    def asc_nulls_last(self) -> "Column":
        return Column("sort", (self, "ascending", "at_end"))

    def desc_nulls_first(self) -> "Column":
        return Column("sort", (self, "descending", "at_start"))

    def desc_nulls_last(self) -> "Column":
        return Column("sort", (self, "descending", "at_end"))

    def _fn(self, fn: str, *params) -> "Column":
        return Column("call", (fn, (self,) + tuple(params)))

    def bitwiseAND(self, other) -> "Column":
        return self._fn("bitwise_and", as_column_or_lit(other))

    def bitwiseOR(self, other) -> "Column":
        return self._fn("bitwise_or", as_column_or_lit(other))

    def bitwiseXOR(self, other) -> "Column":
        return self._fn("bitwise_xor", as_column_or_lit(other))

    def contains(self, other) -> "Column":
        return self._fn("contains", as_column_or_lit(other))

    def startswith(self, other) -> "Column":
        return self._fn("startswith", as_column_or_lit(other))

    def endswith(self, other) -> "Column":
        return self._fn("endswith", as_column_or_lit(other))

    def like(self, pattern: str) -> "Column":
        return self._fn("like", pattern)

    def ilike(self, pattern: str) -> "Column":
        return self._fn("ilike", pattern)

    def rlike(self, pattern: str) -> "Column":
        return self._fn("rlike", pattern)

    def substr(self, startPos, length) -> "Column":
        return self._fn("substring", startPos, length)

    def eqNullSafe(self, other) -> "Column":
        return self._fn("eq_null_safe", as_column_or_lit(other))

    def getItem(self, key) -> "Column":
        return Column("call", ("get_item", (self, as_column_or_lit(key))), name=f"{self._output_name()}[{key}]")

    def getField(self, name: str) -> "Column":
        return Column("call", ("get_field", (self, name)), name=f"{self._output_name()}.{name}")

    def __getitem__(self, key) -> "Column":
        if isinstance(key, slice):
            return self.substr(key.start, key.stop)
        return self.getItem(key)

    def __iter__(self):
        raise TypeError("Column is not iterable")

    def withField(self, fieldName: str, col: "Column") -> "Column":
        return Column("call", ("with_field", (self, fieldName, col)))

    def dropFields(self, *fieldNames: str) -> "Column":
        return Column("call", ("drop_fields", (self, tuple(fieldNames))))

    def __rmod__(self, o): return self._bin("%", o, True)
    def __rpow__(self, o): return self._bin("**", o, True)
    def __rand__(self, o): return self._bin("&", o, True)
    def __ror__(self, o): return self._bin("|", o, True)

    def over(self, window) -> "Column":
        return Column("window", (self, window), self._name)

    def __repr__(self):
        return f"Column<'{self._output_name()}'>"

    # ------------------------------------------------------------------ evaluation
    def references(self) -> List[str]:
        out: List[str] = []

        def walk(node):
            if isinstance(node, Column):
                if node._kind == "col":
                    out.append(node._args[0])
                for a in node._args:
                    walk(a)
            elif isinstance(node, (tuple, list)):
                for a in node:
                    walk(a)
        walk(self)
        return out

    def evaluate(self, table: pa.Table, ctx: Optional[EvalContext] = None):
        """Return an Arrow array (or scalar for pure literals) for this partition."""
        ctx = ctx or EvalContext()
        k, a = self._kind, self._args
        n = table.num_rows
        if k == "col":
            if a[0] not in table.column_names:
                # "alias.column" (the bare column unless a join kept the alias-qualified copy) or "struct.field"
                if "." in a[0]:
                    head, rest = a[0].split(".", 1)
                    if head in table.column_names and pa.types.is_struct(table.schema.field(head).type):
                        value = table.column(head)
                        for part in rest.split("."):
                            value = pc.struct_field(value, part)
                        return value
                    if rest in table.column_names:
                        return table.column(rest)
                raise KeyError(f"cannot resolve column '{a[0]}' given input columns {table.column_names}")
            return table.column(a[0])
        if k == "lit":
            if isinstance(a[0], (list, tuple)):
                one = pa.array([list(a[0])])
                return pa.array([list(a[0])] * n, type=one.type)
            if isinstance(a[0], dict):
                arr = pa.array([list(a[0].items())] * n)
                return pa.MapArray.from_arrays(arr.offsets, pc.struct_field(arr.values, 0), pc.struct_field(arr.values, 1)) \
                    if n else pa.array([], type=pa.map_(pa.string(), pa.string()))
            return pa.scalar(a[0])
        if k == "binop":
            return _binop(a[0], a[1].evaluate(table, ctx), a[2].evaluate(table, ctx))
        if k == "cast":
            return _cast(a[0].evaluate(table, ctx), a[1].arrow, n)
        if k == "isin":
            # SQL three-valued IN: NULL IN (...) is NULL; a miss against a list that holds NULL is NULL
            # (so NOT IN over such a list selects nothing); only a real match is TRUE
            value = _full(a[0].evaluate(table, ctx), n)
            items = list(a[1])
            present = [v for v in items if v is not None]
            hit = pc.is_in(value, value_set=pa.array(present, type=value.type if not present else None)) \
                if present else pa.array(np.zeros(n, dtype=bool))
            if len(present) != len(items):
                hit = pc.if_else(hit, pa.scalar(True), pa.scalar(None, pa.bool_()))
            return pc.if_else(pc.is_valid(value), hit, pa.scalar(None, pa.bool_()))
        if k == "call":
            return _call(a[0], [x.evaluate(table, ctx) if isinstance(x, Column) else x for x in a[1]],
                         n, ctx)
        if k == "udf":
            func, rtype, arg_cols = a
            arrays = [c.evaluate(table, ctx) for c in arg_cols]
            cols = [(_to_array(x, n) if _is_arrayish(x) else None, x) for x in arrays]
            pylists = [c.to_pylist() if c is not None else [s.as_py()] * n for c, s in cols]
            out = [func(*row) for row in zip(*pylists)] if pylists else [func() for _ in range(n)]
            try:
                return pa.array(out, type=rtype.arrow)
            except (pa.ArrowTypeError, pa.ArrowInvalid):
                # the function returned something other than the declared type: text is rendered, anything
                # else goes through a cast of what Arrow infers
                if pa.types.is_string(rtype.arrow):
                    return pa.array([None if v is None else str(v) for v in out], type=pa.string())
                return pc.cast(pa.array(out), rtype.arrow, safe=False)
        if k == "vudf":  # vectorised: func(numpy arrays...) -> numpy array
            func, rtype, arg_cols = a
            arrays = [c.evaluate(table, ctx) for c in arg_cols]
            nps = [np.asarray(_to_array(x, n)) if _is_arrayish(x) else np.full(n, x.as_py())
                   for x in arrays]
            return pa.array(np.asarray(func(*nps)), type=rtype.arrow, from_pandas=True)
        if k == "sort":
            return a[0].evaluate(table, ctx)
        if k == "arrowfn":      # func(*arrow arrays) -> arrow array: whole-partition kernels (sql/ml stages)
            func, arg_cols = a
            arrays = [_full(c.evaluate(table, ctx), n) for c in arg_cols]
            arrays = [x.combine_chunks() if isinstance(x, pa.ChunkedArray) else x for x in arrays]
            return func(*arrays)
        if k == "hof":
            from raydp_b200.sql.fnlib import eval_higher_order
            name, array_cols, body, variables, extra = a
            return eval_higher_order(name, [c.evaluate(table, ctx) for c in array_cols], body, list(variables),
                                     dict(extra), table, ctx)
        if k == "when":
            branches, otherwise = a
            result = otherwise.evaluate(table, ctx) if otherwise is not None else pa.scalar(None)
            for cond, value in reversed(branches):
                # a NULL condition is "not matched": the next branch (or ELSE) decides, as in Spark
                taken = pc.fill_null(_full(cond.evaluate(table, ctx), n), False)
                result = pc.if_else(taken, value.evaluate(table, ctx), result)
            return result
        raise NotImplementedError(f"expression kind {k!r} cannot be evaluated per partition")

    def evaluate_array(self, table: pa.Table, ctx: Optional[EvalContext] = None):
        out = self.evaluate(table, ctx)
        return _full(out, table.num_rows)

    def when(self, condition: "Column", value) -> "Column":
        if self._kind != "when":
            raise TypeError("when() can only be applied on a Column previously generated by when()")
        value = value if isinstance(value, Column) else Column("lit", (value,))
        return Column("when", (self._args[0] + ((condition, value),), self._args[1]))

    def otherwise(self, value) -> "Column":
        value = value if isinstance(value, Column) else Column("lit", (value,))
        return Column("when", (self._args[0], value))


def _cast(value, target: pa.DataType, n: int):
    """Spark SQL cast rules where they differ from Arrow's: timestamps convert to and from numbers as
    SECONDS since the epoch, strings become dates / timestamps leniently, booleans print lower-case."""
    src = value.type
    if pa.types.is_timestamp(src) and (pa.types.is_integer(target) or pa.types.is_floating(target)):
        micros = pc.cast(pc.cast(value, pa.timestamp("us")), pa.int64())
        if pa.types.is_integer(target):
            secs = pc.floor(pc.divide(pc.cast(micros, pa.float64()), 1e6))
            return pc.cast(secs, target, safe=False)
        return pc.cast(pc.divide(pc.cast(micros, pa.float64()), 1e6), target)
    if pa.types.is_timestamp(target) and (pa.types.is_integer(src) or pa.types.is_floating(src)):
        micros = pc.cast(pc.multiply(pc.cast(value, pa.float64()), 1e6), pa.int64(), safe=False)
        return pc.cast(micros, target)
    if (pa.types.is_string(src) or pa.types.is_large_string(src)) and \
            (pa.types.is_timestamp(target) or pa.types.is_date(target)):
        from raydp_b200.sql.fnlib import _as_date, _as_ts
        arr = _full(value, n)
        return _as_date(arr, n, None) if pa.types.is_date(target) else pc.cast(_as_ts(arr, n, EvalContext()), target)
    if pa.types.is_string(target) and pa.types.is_floating(src):
        arr = _full(value, n)
        arr = arr.combine_chunks() if isinstance(arr, pa.ChunkedArray) else arr
        return pa.array([None if v is None else _float_text(v) for v in arr.to_pylist()], type=pa.string())
    if pa.types.is_string(target) and pa.types.is_boolean(src):
        return pc.if_else(value, "true", "false")
    if pa.types.is_string(target) and pa.types.is_timestamp(src):
        from raydp_b200.sql.fnlib import _strftime
        return _strftime(_full(value, n), "%Y-%m-%d %H:%M:%S")
    if (pa.types.is_string(src) or pa.types.is_large_string(src)) and \
            (pa.types.is_integer(target) or pa.types.is_floating(target)):
        # unparsable text becomes null instead of failing the query
        arr = _full(value, n)
        arr = arr.combine_chunks() if isinstance(arr, pa.ChunkedArray) else arr
        trimmed = pc.utf8_trim_whitespace(arr)
        numeric = pc.match_substring_regex(trimmed, r"^[+-]?(\d+\.?\d*([eE][+-]?\d+)?|\.\d+([eE][+-]?\d+)?|NaN|Infinity|-Infinity)$")
        cleaned = pc.if_else(pc.fill_null(numeric, False), trimmed, pa.nulls(len(arr), pa.string()))
        as_float = pc.cast(cleaned, pa.float64())
        return pc.cast(pc.trunc(as_float), target, safe=False) if pa.types.is_integer(target) else \
            pc.cast(as_float, target)
    return pc.cast(value, target, safe=False)


def _float_text(v: float) -> str:
    if v != v:
        return "NaN"
    if v in (float("inf"), float("-inf")):
        return "Infinity" if v > 0 else "-Infinity"
    text = repr(float(v))
    if "e" in text:                                   # 1e+20 -> 1.0E20
        mantissa, exp = text.split("e")
        mantissa = mantissa if "." in mantissa else mantissa + ".0"
        return f"{mantissa}E{int(exp)}"
    return text


def _full(value, n: int):
    """Broadcast scalars to a length-``n`` array."""
    if _is_arrayish(value):
        return value
    if isinstance(value, pa.Scalar):
        if not value.is_valid and pa.types.is_null(value.type):
            return pa.nulls(n)
        return pa.repeat(value, n)
    return pa.repeat(pa.scalar(value), n)


_ARITH = {"+": "add", "-": "subtract", "*": "multiply", "**": "power"}
_CMP = {"<": "less", "<=": "less_equal", ">": "greater", ">=": "greater_equal",
        "==": "equal", "!=": "not_equal"}


def _coerce_pair(left, right):
    """Spark's implicit casts for a comparison: text against a date / timestamp is parsed, text against a
    number becomes a double."""
    lt, rt = left.type, right.type
    is_text = lambda t: pa.types.is_string(t) or pa.types.is_large_string(t)  # noqa: E731
    is_time = lambda t: pa.types.is_date(t) or pa.types.is_timestamp(t)  # noqa: E731
    is_num = lambda t: pa.types.is_integer(t) or pa.types.is_floating(t) or pa.types.is_decimal(t)  # noqa: E731
    if is_text(lt) and is_time(rt):
        return _cast(left, rt, len(left) if _is_arrayish(left) else 1), right
    if is_time(lt) and is_text(rt):
        return left, _cast(right, lt, len(right) if _is_arrayish(right) else 1)
    if pa.types.is_date(lt) and pa.types.is_timestamp(rt):
        return pc.cast(left, rt), right
    if pa.types.is_timestamp(lt) and pa.types.is_date(rt):
        return left, pc.cast(right, lt)
    if is_text(lt) and is_num(rt):
        return _cast(left, pa.float64(), len(left) if _is_arrayish(left) else 1), pc.cast(right, pa.float64())
    if is_num(lt) and is_text(rt):
        return pc.cast(left, pa.float64()), _cast(right, pa.float64(), len(right) if _is_arrayish(right) else 1)
    return left, right


def _scalar_array(value):
    """pa.Scalar -> one-element array (the cast helpers work on arrays)."""
    return pa.array([value.as_py()], type=value.type) if isinstance(value, pa.Scalar) else value


def _binop(op: str, left, right):
    if op in _ARITH:
        return getattr(pc, _ARITH[op])(left, right)
    if op == "/":
        # Spark's `/` is always floating-point division; x / 0 is NULL, not infinity
        num, den = pc.cast(left, pa.float64()), pc.cast(right, pa.float64())
        zero = pc.equal(den, 0.0)
        if isinstance(den, pa.Scalar):
            if den.is_valid and den.as_py() == 0.0:
                return pc.multiply(num, pa.scalar(None, pa.float64()))
            return pc.divide(num, den)
        return pc.divide(num, pc.if_else(pc.fill_null(zero, False), pa.scalar(None, pa.float64()), den))
    if op == "%":
        l_np, r_np = _np(left), _np(right)
        with np.errstate(all="ignore"):
            out = np.fmod(l_np, r_np)
        if not isinstance(out, np.ndarray):
            return pa.scalar(None if r_np == 0 else out)
        mask = np.broadcast_to(np.asarray(r_np) == 0, out.shape)       # x % 0 is NULL
        for side in (left, right):
            if _is_arrayish(side) and side.null_count:
                mask = mask | np.asarray(pc.is_null(side))
        return pa.array(out, mask=mask) if mask.any() else pa.array(out)
    if op in _CMP:
        if left.type != right.type:
            scalar_l, scalar_r = isinstance(left, pa.Scalar), isinstance(right, pa.Scalar)
            a, b = _coerce_pair(_scalar_array(left), _scalar_array(right))
            left = a[0] if scalar_l and _is_arrayish(a) else a
            right = b[0] if scalar_r and _is_arrayish(b) else b
        return getattr(pc, _CMP[op])(left, right)
    if op == "&":
        return pc.and_kleene(left, right)
    if op == "|":
        return pc.or_kleene(left, right)
    raise NotImplementedError(op)


def _np(value):
    if _is_arrayish(value):
        return value.to_numpy(zero_copy_only=False) if isinstance(value, pa.Array) else \
            value.combine_chunks().to_numpy(zero_copy_only=False)
    return value.as_py() if isinstance(value, pa.Scalar) else value


def _ts(value, ctx: EvalContext):
    """Normalise timestamps/strings to timestamp[us] in the session time zone (UTC default)."""
    if _is_arrayish(value):
        t = value.type
    else:
        t = value.type
    if pa.types.is_string(t) or pa.types.is_large_string(t):
        value = parse_timestamp_strings(value)
    elif pa.types.is_date(t):
        value = pc.cast(value, pa.timestamp("us"))
    if pa.types.is_timestamp(value.type) and value.type.tz is not None:
        value = pc.cast(value, pa.timestamp("us"))  # stored as UTC instants
    return value


def parse_timestamp_strings(value):
    """Parse ``2011-11-28 23:11:35 UTC`` / ISO strings (reference ``examples/fake_nyctaxi.csv``)."""
    cleaned = pc.replace_substring_regex(value, r"\s*(UTC|Z)$", "")
    cleaned = pc.replace_substring_regex(cleaned, r"T", " ")
    cleaned = pc.replace_substring_regex(cleaned, r"\.\d+$", "")
    full = pc.strptime(cleaned, format="%Y-%m-%d %H:%M:%S", unit="us", error_is_null=True)
    if (full.is_valid if isinstance(full, pa.Scalar) else full.null_count == 0):
        return full
    day = pc.strptime(pc.utf8_trim_whitespace(cleaned), format="%Y-%m-%d", unit="us", error_is_null=True)
    minute = pc.strptime(cleaned, format="%Y-%m-%d %H:%M", unit="us", error_is_null=True)
    return pc.coalesce(full, day, minute)


def _call(fn: str, args: list, n: int, ctx: EvalContext):
    x = args[0] if args else None
    if fn == "abs":
        return pc.abs(x)
    if fn == "negate":
        return pc.negate(x)
    if fn == "not":
        return pc.invert(x)
    if fn == "is_null":
        return pc.is_null(x)
    if fn == "is_valid":
        return pc.is_valid(x)
    if fn == "log":
        if len(args) == 2:  # log(base, x)
            return pc.divide(pc.ln(pc.cast(args[1], pa.float64())), pc.ln(pc.cast(x, pa.float64())))
        return pc.ln(pc.cast(x, pa.float64()))
    if fn in ("floor", "ceil"):                         # Spark: floor / ceil of a number is a bigint
        return pc.cast(getattr(pc, fn)(pc.cast(x, pa.float64())), pa.int64(), safe=False)
    if fn in ("log1p", "exp", "sqrt", "sin", "cos", "log2", "log10"):
        return getattr(pc, fn)(pc.cast(x, pa.float64()))
    if fn == "round":
        # Spark rounds HALF_UP (2.5 -> 3, -2.5 -> -3); bround() is the half-even one
        digits = args[1] if len(args) > 1 else 0
        digits = digits.as_py() if isinstance(digits, pa.Scalar) else digits
        return pc.round(x, ndigits=int(digits or 0), round_mode="half_towards_infinity")
    if fn == "pow":
        return pc.power(pc.cast(x, pa.float64()), args[1])
    if fn == "rand":
        seed = args[0] if args and args[0] is not None else 0x5eed
        seed = seed.as_py() if isinstance(seed, pa.Scalar) else seed
        rng = np.random.default_rng([int(seed) & 0x7FFFFFFF, ctx.partition_index])
        return pa.array(rng.random(n))
    if fn == "randn":
        seed = args[0] if args and args[0] is not None else 0x5eed
        seed = seed.as_py() if isinstance(seed, pa.Scalar) else seed
        rng = np.random.default_rng([int(seed) & 0x7FFFFFFF, ctx.partition_index, 1])
        return pa.array(rng.standard_normal(n))
    if fn == "monotonically_increasing_id":
        base = ctx.partition_index << 33
        return pa.array(np.arange(base, base + n, dtype=np.int64))
    if fn == "spark_partition_id":
        return pa.array(np.full(n, ctx.partition_index, dtype=np.int32))
    if fn in ("hour", "minute", "second", "month", "quarter", "year"):
        return pc.cast(getattr(pc, fn)(_ts(x, ctx)), pa.int32())
    if fn == "dayofmonth":
        return pc.cast(pc.day(_ts(x, ctx)), pa.int32())
    if fn == "dayofyear":
        return pc.cast(pc.day_of_year(_ts(x, ctx)), pa.int32())
    if fn == "dayofweek":  # Spark: 1 = Sunday ... 7 = Saturday
        return pc.cast(pc.day_of_week(_ts(x, ctx), count_from_zero=False, week_start=7), pa.int32())
    if fn == "weekofyear":
        return pc.cast(pc.iso_week(_ts(x, ctx)), pa.int32())
    if fn == "to_timestamp":
        return _ts(x, ctx)
    if fn == "unix_timestamp":
        return pc.cast(pc.divide(pc.cast(_ts(x, ctx), pa.int64()), 1000000), pa.int64())
    if fn == "coalesce":
        return pc.coalesce(*[_full(a, n) for a in args])
    if fn == "concat":
        return pc.binary_join_element_wise(*[pc.cast(_full(a, n), pa.string()) for a in args], "")
    if fn in ("upper", "lower"):
        return getattr(pc, "utf8_" + fn)(x)
    if fn == "length":
        return pc.utf8_length(x)
    if fn == "greatest":
        return pc.max_element_wise(*args)
    if fn == "least":
        return pc.min_element_wise(*args)
    if fn == "hash":
        # deterministic 32-bit mix of the argument columns (feature crossing helper)
        acc = np.full(n, 42, dtype=np.uint64)
        for a in args:
            v = _np(_full(a, n))
            if v.dtype.kind in "OUS":
                v = np.array([hash_str(s) for s in v], dtype=np.uint64)
            else:
                v = v.astype(np.int64, copy=False).view(np.uint64)
            acc = (acc ^ v) * np.uint64(0x9E3779B97F4A7C15)
            acc ^= acc >> np.uint64(29)
        return pa.array((acc & np.uint64(0x7FFFFFFF)).astype(np.int32))
    if fn == "array":
        # element j of row i = args[j][i]; done in Arrow so strings and nulls survive (numpy stacking
        # would turn them into objects)
        if not args:                                     # array(): an empty array per row
            return pa.array([[]] * n, type=pa.list_(pa.null()))
        parts = []
        for a in args:
            v = _full(a, n)
            v = v.combine_chunks() if isinstance(v, pa.ChunkedArray) else (
                v if isinstance(v, pa.Array) else pa.array(v))
            parts.append(v)
        common = parts[0].type
        for v in parts[1:]:
            if v.type != common:
                common = pa.float64() if (pa.types.is_floating(common) or pa.types.is_floating(v.type)) else (
                    pa.int64() if pa.types.is_integer(common) and pa.types.is_integer(v.type) else pa.string())
        parts = [v.cast(common) for v in parts]
        k = len(parts)
        flat = pa.concat_arrays(parts)                                   # column-major: [col0 | col1 | ...]
        order = (np.arange(n, dtype=np.int64)[:, None] + np.arange(k, dtype=np.int64)[None, :] * n).reshape(-1)
        values = flat.take(pa.array(order))                              # row-major
        offsets = pa.array(np.arange(0, (n + 1) * k, k, dtype=np.int32))
        return pa.ListArray.from_arrays(offsets, values)
    from raydp_b200.sql.fnlib import SCALAR
    impl = SCALAR.get(fn)
    if impl is not None:
        return impl(args, n, ctx)
    raise NotImplementedError(f"SQL function {fn!r} is not supported by the built-in engine")


def hash_str(s) -> int:
    h = 1469598103934665603
    for ch in str(s).encode():
        h = ((h ^ ch) * 1099511628211) & 0xFFFFFFFFFFFFFFFF
    return h


def as_column_or_lit(value) -> Column:
    """A Column stays; anything else (strings included) becomes a literal."""
    return value if isinstance(value, Column) else Column("lit", (value,))


def as_column(value) -> Column:
    if isinstance(value, Column):
        return value
    if isinstance(value, str):
        return Column("col", (value,))
    return Column("lit", (value,))
