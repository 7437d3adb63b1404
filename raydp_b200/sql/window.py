"""Window specifications (``Window.partitionBy(...).orderBy(...).rowsBetween(...)``).

Parity: the reference's DLRM notebook uses ``row_number().over(Window.partitionBy().orderBy(desc(...)))``
(cell 6, frequency ranking of categorical values); user code written against pyspark uses the rest
of the surface -- ranking functions, ``lag`` / ``lead``, aggregates over ROWS / RANGE frames.  The
evaluation lives in :mod:`raydp_b200.sql.windowing`.
"""
from __future__ import annotations

import sys

from raydp_b200.sql.column import as_column


class WindowSpec:
    def __init__(self, partition_by=(), order_by=(), frame=None):
        self._partition_by = tuple(partition_by)
        self._order_by = tuple(order_by)
        self._frame = frame                       # None | ("rows"|"range", start, end)

    def partitionBy(self, *cols) -> "WindowSpec":
        cols = cols[0] if len(cols) == 1 and isinstance(cols[0], (list, tuple)) else cols
        return WindowSpec(tuple(as_column(c) for c in cols), self._order_by, self._frame)

    def orderBy(self, *cols) -> "WindowSpec":
        cols = cols[0] if len(cols) == 1 and isinstance(cols[0], (list, tuple)) else cols
        return WindowSpec(self._partition_by, tuple(as_column(c) for c in cols), self._frame)

    def rowsBetween(self, start: int, end: int) -> "WindowSpec":
        return WindowSpec(self._partition_by, self._order_by, ("rows", _clamp(start), _clamp(end)))

    def rangeBetween(self, start: int, end: int) -> "WindowSpec":
        return WindowSpec(self._partition_by, self._order_by, ("range", _clamp(start), _clamp(end)))


def _clamp(bound: int) -> int:
    # pyspark treats anything at or beyond +-2^63 (and its own sentinels) as unbounded
    if bound <= Window.unboundedPreceding or bound <= -(1 << 62):
        return Window.unboundedPreceding
    if bound >= Window.unboundedFollowing or bound >= (1 << 62):
        return Window.unboundedFollowing
    return int(bound)


class Window:
    unboundedPreceding: int = -sys.maxsize
    unboundedFollowing: int = sys.maxsize
    currentRow: int = 0

    @staticmethod
    def partitionBy(*cols) -> WindowSpec:
        return WindowSpec().partitionBy(*cols)

    @staticmethod
    def orderBy(*cols) -> WindowSpec:
        return WindowSpec().orderBy(*cols)

    @staticmethod
    def rowsBetween(start: int, end: int) -> WindowSpec:
        return WindowSpec().rowsBetween(start, end)

    @staticmethod
    def rangeBetween(start: int, end: int) -> WindowSpec:
        return WindowSpec().rangeBetween(start, end)
