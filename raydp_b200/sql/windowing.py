"""Ordering and window functions.

Window expressions (``F.rank().over(Window.partitionBy("k").orderBy("v"))``, running sums, ``lag`` /
``lead``, frames) are what pyspark users reach for right after ``groupBy``; the reference's DLRM
notebook ranks categorical values with ``row_number`` (cell 6).  A window needs every row of a
partition-key group in one place, so the frame is gathered, sorted once by (partition keys, order
keys) and every function is a numpy pass over the sorted arrays: segment starts for the partitions,
peer-group boundaries for ties, prefix sums for running aggregates.

Sorting follows Spark: ascending puts nulls first, descending puts them last, unless the key says
otherwise (``asc_nulls_last()`` ...).
"""
from __future__ import annotations

import sys
from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

from raydp_b200.sql.column import Column, EvalContext, as_column, _full

UNBOUNDED_PRECEDING = -sys.maxsize
UNBOUNDED_FOLLOWING = sys.maxsize
CURRENT_ROW = 0


# ------------------------------------------------------------------------------------------ ordering
def sort_spec(col, ascending: bool = True) -> Tuple[Column, str, str]:
    """``(expression, "ascending"|"descending", "at_start"|"at_end")`` for one ORDER BY item."""
    if isinstance(col, Column) and col._kind == "sort":
        direction = col._args[1]
        placement = col._args[2] if len(col._args) > 2 else ("at_start" if direction == "ascending" else "at_end")
        return col._args[0], direction, placement
    direction = "ascending" if ascending else "descending"
    return as_column(col), direction, "at_start" if ascending else "at_end"


def sort_indices(table: pa.Table, specs: Sequence[Tuple[Column, str, str]], ctx: Optional[EvalContext] = None):
    """Stable sort permutation of ``table`` for the given ORDER BY items (expressions allowed)."""
    if not specs or table.num_rows == 0:
        return pa.array(np.arange(table.num_rows, dtype=np.int64))
    work: Dict[str, Any] = {}
    keys: List[Tuple[str, str]] = []
    for i, (expr, direction, placement) in enumerate(specs):
        values = _full(expr.evaluate(table, ctx), table.num_rows)
        if pa.types.is_null(values.type):
            continue
        if isinstance(values, pa.ChunkedArray):
            values = values.combine_chunks()
        if values.null_count:
            # a leading indicator key puts this column's nulls where asked, independent of the others
            work[f"n{i}"] = pc.cast(pc.is_null(values), pa.int8())
            keys.append((f"n{i}", "descending" if placement == "at_start" else "ascending"))
        work[f"k{i}"] = values
        keys.append((f"k{i}", direction))
    if not keys:
        return pa.array(np.arange(table.num_rows, dtype=np.int64))
    return pc.sort_indices(pa.table(work), sort_keys=keys)


def sort_table(table: pa.Table, cols, ascending=True, ctx: Optional[EvalContext] = None) -> pa.Table:
    if len(cols) == 1 and isinstance(cols[0], (list, tuple)):
        cols = tuple(cols[0])
    asc = ascending if isinstance(ascending, (list, tuple)) else [ascending] * len(cols)
    specs = [sort_spec(c, bool(a)) for c, a in zip(cols, asc)]
    return table.take(sort_indices(table, specs, ctx))


# ------------------------------------------------------------------------------------------ windows
def _change_points(arrays: List[pa.Array], n: int) -> np.ndarray:
    """``change[i]`` is True where row ``i`` differs from row ``i-1`` in any of ``arrays`` (nulls equal)."""
    change = np.zeros(n, dtype=bool)
    if n:
        change[0] = True
    for v in arrays:
        if isinstance(v, pa.ChunkedArray):
            v = v.combine_chunks()
        if n < 2:
            continue
        a, b = v.slice(1), v.slice(0, n - 1)
        neq = pc.fill_null(pc.not_equal(a, b), False)
        null_diff = pc.xor(pc.is_null(a), pc.is_null(b))
        change[1:] |= np.asarray(pc.or_(neq, null_diff)).astype(bool)
    return change


def _starts_ends(change: np.ndarray) -> Tuple[np.ndarray, np.ndarray]:
    """Per row: index of the first row of its group and one past the last."""
    n = len(change)
    idx = np.arange(n)
    starts = np.maximum.accumulate(np.where(change, idx, 0))
    boundaries = np.flatnonzero(change)
    ends_of_group = np.append(boundaries[1:], n)
    group_id = np.cumsum(change) - 1
    return starts, ends_of_group[group_id] if n else np.zeros(0, dtype=np.int64)


class _Frame:
    """The sorted table plus partition / peer structure shared by every function over one window spec."""

    def __init__(self, table: pa.Table, spec, ctx: EvalContext):
        n = table.num_rows
        part_specs = [(c, "ascending", "at_start") for c in spec._partition_by]
        order_specs = [sort_spec(c) for c in spec._order_by]
        idx = sort_indices(table, part_specs + order_specs, ctx)
        self.table = table.take(idx)
        self.n = n
        self.ctx = ctx
        self.spec = spec
        part_arrays = [_full(c.evaluate(self.table, ctx), n) for c in spec._partition_by]
        self.order_arrays = [_full(s[0].evaluate(self.table, ctx), n) for s in order_specs]
        self.order_specs = order_specs
        self.part_change = _change_points(part_arrays, n)
        self.peer_change = self.part_change | _change_points(self.order_arrays, n) if order_specs else \
            self.part_change.copy()
        self.seg_start, self.seg_end = _starts_ends(self.part_change)
        self.peer_start, self.peer_end = _starts_ends(self.peer_change)
        self.idx = np.arange(n)

    # -- frame bounds ------------------------------------------------------------------------------
    def bounds(self) -> Tuple[np.ndarray, np.ndarray]:
        """Inclusive ``[lo, hi]`` row range (in sorted order) of the frame of every row."""
        spec = self.spec
        kind, lo_off, hi_off = spec._frame if spec._frame is not None else (
            ("range", UNBOUNDED_PRECEDING, CURRENT_ROW) if spec._order_by else
            ("rows", UNBOUNDED_PRECEDING, UNBOUNDED_FOLLOWING))
        first, last = self.seg_start, self.seg_end - 1
        if kind == "rows":
            lo = first if lo_off <= UNBOUNDED_PRECEDING else np.maximum(first, self.idx + lo_off)
            hi = last if hi_off >= UNBOUNDED_FOLLOWING else np.minimum(last, self.idx + hi_off)
            return lo, hi
        # RANGE: offsets are distances in the (single, numeric) order key; 0 means "my peers"
        lo = first if lo_off <= UNBOUNDED_PRECEDING else (self.peer_start if lo_off == 0 else None)
        hi = last if hi_off >= UNBOUNDED_FOLLOWING else (self.peer_end - 1 if hi_off == 0 else None)
        if lo is None or hi is None:
            if len(self.order_arrays) != 1:
                raise ValueError("a RANGE frame with offsets needs exactly one ORDER BY expression")
            key = np.asarray(pc.cast(self.order_arrays[0], pa.float64()).to_numpy(zero_copy_only=False),
                             dtype=np.float64)
            if self.order_specs[0][1] == "descending":
                key = -key
            lo_out, hi_out = (lo if lo is not None else np.empty(self.n, dtype=np.int64)), \
                (hi if hi is not None else np.empty(self.n, dtype=np.int64))
            for s in np.flatnonzero(self.part_change):
                e = self.seg_end[s]
                seg = key[s:e]
                if lo is None:
                    lo_out[s:e] = s + np.searchsorted(seg, seg + lo_off, side="left")
                if hi is None:
                    hi_out[s:e] = s + np.searchsorted(seg, seg + hi_off, side="right") - 1
            lo, hi = lo_out, hi_out
        return lo, hi

    # -- ranking -----------------------------------------------------------------------------------
    def ranking(self, fn: str, params: tuple) -> pa.Array:
        n = self.n
        pos = self.idx - self.seg_start                         # 0-based position in the partition
        size = self.seg_end - self.seg_start
        if fn == "row_number":
            return pa.array((pos + 1).astype(np.int32))
        if fn == "rank":
            return pa.array((self.peer_start - self.seg_start + 1).astype(np.int32))
        if fn == "dense_rank":
            peers_seen = np.cumsum(self.peer_change)
            return pa.array((peers_seen - peers_seen[self.seg_start] + 1).astype(np.int32))
        if fn == "percent_rank":
            rank0 = (self.peer_start - self.seg_start).astype(np.float64)
            return pa.array(np.where(size > 1, rank0 / np.maximum(size - 1, 1), 0.0))
        if fn == "cume_dist":
            return pa.array((self.peer_end - self.seg_start) / size.astype(np.float64))
        if fn == "ntile":
            buckets = int(params[0])
            base, extra = size // buckets, size % buckets
            cut = extra * (base + 1)                            # rows that sit in the larger buckets
            tile = np.where(pos < cut, pos // np.maximum(base + 1, 1),
                            extra + (pos - cut) // np.maximum(base, 1))
            return pa.array((tile + 1).astype(np.int32))
        if fn in ("lag", "lead"):
            target, offset, default = params
            offset = int(offset) * (-1 if fn == "lag" else 1)
            values = _full(target.evaluate(self.table, self.ctx), n)
            values = values.combine_chunks() if isinstance(values, pa.ChunkedArray) else values
            src = self.idx + offset
            inside = (src >= self.seg_start) & (src < self.seg_end)
            taken = values.take(pa.array(np.where(inside, src, 0))) if n else values
            fill = pa.nulls(n, values.type) if default is None else \
                pc.cast(_full(as_column_or_lit(default).evaluate(self.table, self.ctx), n), values.type)
            return pc.if_else(pa.array(inside), taken, fill)
        raise NotImplementedError(f"window function {fn!r}")

    # -- aggregates over the frame -----------------------------------------------------------------
    def aggregate(self, fn: str, target: Optional[Column], params: tuple = ()) -> pa.Array:
        n = self.n
        lo, hi = self.bounds()
        empty = hi < lo
        lo_c, hi_c = np.minimum(lo, np.maximum(n - 1, 0)), np.maximum(hi, 0)
        if target is None:                                      # count(*)
            return pa.array(np.where(empty, 0, hi - lo + 1).astype(np.int64))
        values = _full(target.evaluate(self.table, self.ctx), n)
        values = values.combine_chunks() if isinstance(values, pa.ChunkedArray) else values
        valid = np.asarray(pc.is_valid(values)).astype(bool)
        numeric = pa.types.is_integer(values.type) or pa.types.is_floating(values.type) or \
            pa.types.is_boolean(values.type) or pa.types.is_decimal(values.type)
        if fn in ("count", "sum", "mean", "stddev", "variance", "stddev_pop", "var_pop") and (numeric or fn == "count"):
            cnt_prefix = np.concatenate([[0], np.cumsum(valid)])
            cnt = np.where(empty, 0, cnt_prefix[hi_c + 1] - cnt_prefix[lo_c])
            if fn == "count":
                return pa.array(cnt.astype(np.int64))
            is_int = pa.types.is_integer(values.type) and fn == "sum"
            x = np.asarray(pc.fill_null(pc.cast(values, pa.int64() if is_int else pa.float64()), 0))
            pre = np.concatenate([[0], np.cumsum(x)])
            total = pre[hi_c + 1] - pre[lo_c]
            none = cnt == 0
            if fn == "sum":
                return pa.array(total, mask=none)
            if fn == "mean":
                return pa.array(total / np.maximum(cnt, 1), mask=none)
            pre2 = np.concatenate([[0.0], np.cumsum(x.astype(np.float64) ** 2)])
            sq = pre2[hi_c + 1] - pre2[lo_c]
            ddof = 0 if fn in ("stddev_pop", "var_pop") else 1
            denom = cnt - ddof
            var = (sq - total * total / np.maximum(cnt, 1)) / np.where(denom > 0, denom, 1)
            var = np.maximum(var, 0.0)
            out = np.sqrt(var) if fn.startswith("stddev") else var
            return pa.array(out, mask=denom <= 0)
        if fn in ("min", "max") and numeric and not pa.types.is_boolean(values.type) and n and \
                np.array_equal(lo, self.seg_start):
            # frames that start at the partition start (running and whole-partition frames): a per-partition
            # cumulative min / max, read at each row's frame end
            import pandas as pd
            x = np.asarray(pc.cast(values, pa.float64()).to_numpy(zero_copy_only=False), dtype=np.float64)
            x = np.where(valid, x, np.inf if fn == "min" else -np.inf)
            seg_id = np.cumsum(self.part_change) - 1
            grouped = pd.Series(x).groupby(seg_id)
            acc = (grouped.cummin() if fn == "min" else grouped.cummax()).to_numpy()
            cnt_prefix = np.concatenate([[0], np.cumsum(valid)])
            none = (cnt_prefix[hi_c + 1] - cnt_prefix[lo_c] == 0) | empty
            out_vals = acc[hi_c]
            result = pa.array(np.where(none, 0.0, out_vals), mask=none)
            return pc.cast(result, values.type) if pa.types.is_integer(values.type) else result
        # everything else: evaluate the frame slice; frames shared by a whole group are computed once
        from raydp_b200.sql.aggregates import reduce_values
        out: List[Any] = [None] * n
        cache: Dict[Tuple[int, int], Any] = {}
        extra = None
        if fn in ("min_by", "max_by", "corr", "covar_pop", "covar_samp") and params:
            extra = _full(params[0].evaluate(self.table, self.ctx), n)
            extra = extra.combine_chunks() if isinstance(extra, pa.ChunkedArray) else extra
        for i in range(n):
            if empty[i]:
                continue
            key = (int(lo[i]), int(hi[i]))
            if key not in cache:
                window = values.slice(key[0], key[1] - key[0] + 1)
                other = extra.slice(key[0], key[1] - key[0] + 1) if extra is not None else None
                cache[key] = reduce_values(fn, window, other, params)
            out[i] = cache[key]
        from raydp_b200.sql.aggregates import result_type
        return pa.array(out, type=result_type(fn, values.type))


def as_column_or_lit(v) -> Column:
    return v if isinstance(v, Column) else Column("lit", (v,))


def evaluate_windows(table: pa.Table, requests: List[Tuple[str, Column]], ctx: Optional[EvalContext] = None) -> pa.Table:
    """Append one column per ``(name, window expression)``; the result is sorted by the last spec."""
    ctx = ctx or EvalContext()
    for name, wcol in requests:
        fn, spec = wcol._args
        frame = _Frame(table, spec, ctx)
        if fn._kind == "rank_fn":
            values = frame.ranking(fn._args[0], tuple(fn._args[1:]))
        elif fn._kind == "agg":
            values = frame.aggregate(fn._args[0], fn._args[1], tuple(fn._args[2:]))
        else:
            raise TypeError(f"{fn!r} is not a window function; use an aggregate or a ranking function with over()")
        table = frame.table
        if name in table.column_names:
            table = table.set_column(table.column_names.index(name), name, values)
        else:
            table = table.append_column(name, values)
    return table


def extract_windows(expr: Column, prefix: str, found: List[Tuple[str, Column]]) -> Column:
    """Replace every window node inside ``expr`` with a reference to a temporary column."""
    if not isinstance(expr, Column):
        return expr
    if expr._kind == "window":
        tmp = f"{prefix}{len(found)}"
        found.append((tmp, expr))
        return Column("col", (tmp,), expr._name)

    def walk(node):
        if isinstance(node, Column):
            return extract_windows(node, prefix, found)
        if isinstance(node, tuple):
            return tuple(walk(x) for x in node)
        if isinstance(node, list):
            return [walk(x) for x in node]
        return node
    if expr._kind in ("col", "lit", "udf_fn"):
        return expr
    return Column(expr._kind, tuple(walk(a) for a in expr._args), expr._name)


def has_window(expr) -> bool:
    if isinstance(expr, Column):
        return expr._kind == "window" or any(has_window(a) for a in expr._args)
    if isinstance(expr, (tuple, list)):
        return any(has_window(a) for a in expr)
    return False
