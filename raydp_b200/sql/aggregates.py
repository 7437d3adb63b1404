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


def _reduce_groups_numeric(fn: str, lists: pa.Array, params: tuple) -> Optional[pa.Array]:
    """Segmented numpy versions of the numeric reducers (one pass over all groups instead of a Python loop
    per group): moments through ``bincount``, order statistics through one (group, value) sort."""
    vtype = lists.type.value_type
    if not (pa.types.is_integer(vtype) or pa.types.is_floating(vtype)) or lists.null_count:
        return None
    if fn not in ("median", "percentile_approx", "stddev_pop", "var_pop", "stddev_samp", "var_samp", "skewness",
                  "kurtosis", "count_distinct", "approx_count_distinct", "sum_distinct", "product"):
        return None
    if fn == "percentile_approx" and isinstance(params[0], (list, tuple)):
        return None
    g = len(lists)
    lengths = np.asarray(pc.list_value_length(lists)).astype(np.int64)
    group = np.repeat(np.arange(g, dtype=np.int64), lengths)
    flat = pc.list_flatten(lists)
    keep = np.asarray(pc.is_valid(flat))
    if pa.types.is_floating(vtype):
        keep &= ~np.asarray(pc.fill_null(pc.is_nan(flat), False)) if fn in ("median", "percentile_approx") else True
    raw = flat.to_numpy(zero_copy_only=False)
    group, vals = group[keep], raw[keep]
    x = vals.astype(np.float64)
    n = np.bincount(group, minlength=g).astype(np.float64)
    empty = n == 0
    if fn in ("median", "percentile_approx"):
        order = np.lexsort((x, group))
        sorted_vals, starts = vals[order], np.concatenate([[0], np.cumsum(n)[:-1]]).astype(np.int64)
        cnt = n.astype(np.int64)
        if fn == "median":
            lo = starts + np.maximum(cnt - 1, 0) // 2
            hi = starts + cnt // 2
            safe_lo, safe_hi = np.minimum(lo, max(len(x) - 1, 0)), np.minimum(hi, max(len(x) - 1, 0))
            out = (sorted_vals[safe_lo].astype(np.float64) + sorted_vals[safe_hi].astype(np.float64)) / 2.0 if len(x) else \
                np.zeros(g)
            return pa.array(out, mask=empty)
        q = float(params[0])
        k = np.maximum(np.ceil(q * cnt).astype(np.int64) - 1, 0)
        idx = np.minimum(starts + k, max(len(x) - 1, 0))
        out = sorted_vals[idx] if len(x) else np.zeros(g, dtype=vals.dtype)
        return pa.array(out, mask=empty)
    if fn in ("count_distinct", "approx_count_distinct", "sum_distinct"):
        pairs = np.unique(np.stack([group.astype(np.float64), x], axis=1), axis=0) if len(x) else np.zeros((0, 2))
        gid = pairs[:, 0].astype(np.int64)
        if fn == "sum_distinct":
            total = np.bincount(gid, weights=pairs[:, 1], minlength=g)
            if pa.types.is_integer(vtype):
                return pa.array(total.astype(np.int64), mask=empty)
            return pa.array(total, mask=empty)
        return pa.array(np.bincount(gid, minlength=g).astype(np.int64))
    if fn == "product":
        out = np.ones(g)
        np.multiply.at(out, group, x)
        return pa.array(out, mask=empty)
    safe_n = np.maximum(n, 1.0)
    mean = np.bincount(group, weights=x, minlength=g) / safe_n
    d = x - mean[group]
    m2 = np.bincount(group, weights=d * d, minlength=g)
    if fn in ("stddev_pop", "var_pop", "stddev_samp", "var_samp"):
        ddof = 0.0 if fn.endswith("_pop") else 1.0
        bad = (n - ddof) <= 0
        var = m2 / np.where(bad, 1.0, n - ddof)
        return pa.array(np.sqrt(var) if fn.startswith("stddev") else var, mask=bad)
    m2 = m2 / safe_n
    if fn == "skewness":
        m3 = np.bincount(group, weights=d ** 3, minlength=g) / safe_n
        bad = empty | (m2 <= 0)
        return pa.array(m3 / np.where(bad, 1.0, m2) ** 1.5, mask=bad)
    m4 = np.bincount(group, weights=d ** 4, minlength=g) / safe_n
    bad = empty | (m2 <= 0)
    return pa.array(m4 / np.where(bad, 1.0, m2 * m2) - 3.0, mask=bad)


def reduce_groups(fn: str, lists: pa.ChunkedArray, params: tuple, input_type: pa.DataType) -> pa.Array:
    """Final stage of a gathered aggregate: one list of values per group -> one value per group."""
    if isinstance(lists, pa.ChunkedArray):
        lists = lists.combine_chunks()
    if fn == "collect_list" and lists.null_count == 0 and lists.values.null_count == 0:
        return lists                                    # already one list per group, nothing to drop
    if fn in ("first_value", "any_value", "first", "last", "last_value") and not (params and params[-1]) \
            and len(lists) and lists.null_count == 0:
        lengths = pc.list_value_length(lists)
        if pc.min(lengths).as_py() > 0:                 # vectorised: first / last element of every group
            offsets = np.asarray(lists.offsets)
            pick = offsets[:-1] if fn in ("first_value", "any_value", "first") else offsets[1:] - 1
            return lists.values.take(pa.array(pick))
    fast = _reduce_groups_numeric(fn, lists, params)
    if fast is not None:
        return fast
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
