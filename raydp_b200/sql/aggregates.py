"""Aggregates beyond the mergeable basics.

``count / sum / min / max / avg / stddev / variance / first / countDistinct`` run as a two-stage
aggregation with small partial states (``sql/executor.py::_partial_agg``).  Everything else --
``collect_list``, ``percentile_approx``, ``corr``, ``skewness``, ``max_by`` ... -- has no compact
partial state worth maintaining here: the first stage gathers each group's values into a list, the
lists are concatenated per group and :func:`reduce_values` computes the answer from the group's full
value array.  The same reducer evaluates aggregates over window frames (``sql/windowing.py``).
"""
from __future__ import annotations

import math
from typing import Any, List, Optional

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

# aggregates with a compact, mergeable state (handled by executor._partial_agg / dataframe._final_agg)
MERGEABLE = {"count", "sum", "min", "max", "mean", "stddev", "variance", "first", "count_distinct"}

# aggregates over two columns: the engine packs them into one struct column {"0": x, "1": y}
PAIRED = {"corr", "covar_pop", "covar_samp", "min_by", "max_by"}


def _drop_nulls(values: pa.Array) -> pa.Array:
    return values.drop_null() if values.null_count else values


def _floats(values: pa.Array) -> np.ndarray:
    return np.asarray(pc.cast(_drop_nulls(values), pa.float64()).to_numpy(zero_copy_only=False), dtype=np.float64)


def _pair(values: pa.Array, other: Optional[pa.Array]):
    if other is None:
        if isinstance(values, pa.ChunkedArray):
            values = values.combine_chunks()
        values, other = pc.struct_field(values, "0"), pc.struct_field(values, "1")
    keep = pc.and_(pc.is_valid(values), pc.is_valid(other))
    return values.filter(keep), other.filter(keep)


def _quantile(sorted_vals: np.ndarray, q: float):
    """Spark's ``percentile_approx`` returns an element of the data: the smallest value whose rank covers q."""
    if len(sorted_vals) == 0:
        return None
    k = max(int(math.ceil(q * len(sorted_vals))) - 1, 0)
    return sorted_vals[k].item()


def reduce_values(fn: str, values, other=None, params: tuple = ()) -> Any:
    """One group's (or one window frame's) value array -> the aggregate, as a Python value."""
    if isinstance(values, pa.ChunkedArray):
        values = values.combine_chunks()
    if fn == "collect_list":
        return _drop_nulls(values).to_pylist()
    if fn == "collect_set":
        return pc.unique(_drop_nulls(values)).to_pylist()
    if fn in ("first", "first_value", "any_value"):
        ignore_nulls = bool(params[-1]) if params else False
        pool = _drop_nulls(values) if ignore_nulls else values
        return pool[0].as_py() if len(pool) else None
    if fn in ("last", "last_value"):
        ignore_nulls = bool(params[-1]) if params else False
        pool = _drop_nulls(values) if ignore_nulls else values
        return pool[-1].as_py() if len(pool) else None
    if fn == "nth_value":
        k = int(params[0])
        return values[k - 1].as_py() if 0 < k <= len(values) else None
    if fn == "count":
        return len(values) - values.null_count
    if fn == "count_if":
        return int(pc.sum(pc.cast(pc.fill_null(values, False), pa.int64())).as_py() or 0)
    if fn in ("min", "max"):
        live = _drop_nulls(values)
        return getattr(pc, fn)(live).as_py() if len(live) else None
    if fn == "sum":
        live = _drop_nulls(values)
        return pc.sum(live).as_py() if len(live) else None
    if fn == "mean":
        live = _drop_nulls(values)
        return pc.mean(live).as_py() if len(live) else None
    if fn == "count_distinct_rows":       # count(DISTINCT a, b): rows with a null in any column do not count
        rows = [tuple(r.values()) for r in values.to_pylist() if r is not None]
        return len({repr(r) for r in rows if all(v is not None for v in r)})
    if fn in ("count_distinct", "approx_count_distinct"):
        return len(pc.unique(_drop_nulls(values)))
    if fn == "sum_distinct":
        live = pc.unique(_drop_nulls(values))
        return pc.sum(live).as_py() if len(live) else None
    if fn in ("stddev", "stddev_samp", "variance", "var_samp", "stddev_pop", "var_pop"):
        x = _floats(values)
        ddof = 0 if fn.endswith("_pop") else 1
        if len(x) - ddof <= 0:
            return None
        var = float(np.var(x, ddof=ddof))
        return math.sqrt(var) if fn.startswith("stddev") else var
    if fn == "skewness":
        x = _floats(values)
        if len(x) == 0:
            return None
        d = x - x.mean()
        m2, m3 = np.mean(d ** 2), np.mean(d ** 3)
        return float(m3 / m2 ** 1.5) if m2 > 0 else None
    if fn == "kurtosis":
        x = _floats(values)
        if len(x) == 0:
            return None
        d = x - x.mean()
        m2, m4 = np.mean(d ** 2), np.mean(d ** 4)
        return float(m4 / (m2 * m2) - 3.0) if m2 > 0 else None
    if fn == "product":
        x = _floats(values)
        return float(np.prod(x)) if len(x) else None
    if fn in ("bool_and", "every"):
        live = _drop_nulls(values)
        return bool(pc.all(live).as_py()) if len(live) else None
    if fn in ("bool_or", "some", "any"):
        live = _drop_nulls(values)
        return bool(pc.any(live).as_py()) if len(live) else None
    if fn in ("bit_and", "bit_or", "bit_xor"):
        live = _drop_nulls(values).to_numpy(zero_copy_only=False)
        if len(live) == 0:
            return None
        op = {"bit_and": np.bitwise_and, "bit_or": np.bitwise_or, "bit_xor": np.bitwise_xor}[fn]
        return int(op.reduce(live.astype(np.int64)))
    if fn == "median":
        x = _floats(values)
        return float(np.median(x)) if len(x) else None
    if fn == "percentile":
        x = _floats(values)
        q = params[0]
        if len(x) == 0:
            return None
        return [float(np.quantile(x, p)) for p in q] if isinstance(q, (list, tuple)) else float(np.quantile(x, q))
    if fn == "percentile_approx":
        live = _drop_nulls(values)
        ordered = np.sort(live.to_numpy(zero_copy_only=False))
        q = params[0]
        return [_quantile(ordered, p) for p in q] if isinstance(q, (list, tuple)) else _quantile(ordered, q)
    if fn == "mode":
        live = _drop_nulls(values)
        if len(live) == 0:
            return None
        counts = pc.value_counts(live)
        best = int(np.argmax(np.asarray(pc.struct_field(counts, "counts"))))
        return pc.struct_field(counts, "values")[best].as_py()
    if fn in ("corr", "covar_pop", "covar_samp"):
        a, b = _pair(values, other)
        x, y = _floats(a), _floats(b)
        if len(x) == 0:
            return None
        if fn == "corr":
            if len(x) < 2 or x.std() == 0 or y.std() == 0:
                return None if len(x) < 2 else float("nan")
            return float(np.corrcoef(x, y)[0, 1])
        ddof = 0 if fn == "covar_pop" else 1
        if len(x) - ddof <= 0:
            return None
        return float(np.sum((x - x.mean()) * (y - y.mean())) / (len(x) - ddof))
    if fn in ("min_by", "max_by"):
        a, b = (values, other) if other is not None else (pc.struct_field(values, "0"), pc.struct_field(values, "1"))
        keep = pc.is_valid(b)
        a, b = a.filter(keep), b.filter(keep)
        if len(b) == 0:
            return None
        order = pc.array_sort_indices(b, order="ascending" if fn == "min_by" else "descending")
        return a[order[0].as_py()].as_py()
    raise NotImplementedError(f"aggregate {fn!r} is not supported by the built-in engine")


def result_type(fn: str, input_type: pa.DataType) -> Optional[pa.DataType]:
    """Arrow type of the aggregate's output column (``None`` lets Arrow infer it)."""
    if fn in ("collect_list", "collect_set"):
        return pa.list_(input_type)
    if fn in ("first", "last", "first_value", "last_value", "nth_value", "any_value", "min", "max", "mode"):
        return input_type
    if fn in ("min_by", "max_by"):
        return input_type.field("0").type if pa.types.is_struct(input_type) else input_type
    if fn in ("count", "count_if", "count_distinct", "count_distinct_rows", "approx_count_distinct", "bit_and", "bit_or", "bit_xor"):
        return pa.int64()
    if fn in ("bool_and", "every", "bool_or", "some", "any"):
        return pa.bool_()
    if fn in ("sum", "sum_distinct"):
        return pa.int64() if pa.types.is_integer(input_type) else pa.float64()
    if fn == "percentile_approx":
        return None
    if fn == "percentile":
        return None
    return pa.float64()


def reduce_groups(fn: str, lists: pa.ChunkedArray, params: tuple, input_type: pa.DataType) -> pa.Array:
    """Final stage of a gathered aggregate: one list of values per group -> one value per group."""
    if isinstance(lists, pa.ChunkedArray):
        lists = lists.combine_chunks()
    out: List[Any] = []
    for i in range(len(lists)):
        cell = lists[i]
        values = cell.values if cell.is_valid and cell.values is not None else pa.array([], type=input_type)
        out.append(reduce_values(fn, values, None, params))
    rtype = result_type(fn, input_type)
    if fn in ("percentile_approx", "percentile"):
        q = params[0] if params else 0.5
        base = input_type if fn == "percentile_approx" else pa.float64()
        rtype = pa.list_(base) if isinstance(q, (list, tuple)) else base
    return pa.array(out, type=rtype)
