"""Window specifications (``Window.partitionBy(...).orderBy(...)``).

Parity: only ``row_number().over(Window.partitionBy().orderBy(desc(...)))`` is used by the
reference (DLRM notebook cell 6, frequency ranking of categorical values).
"""
from __future__ import annotations

from raydp_b200.sql.column import as_column


class WindowSpec:
    def __init__(self, partition_by=(), order_by=()):
        self._partition_by = tuple(partition_by)
        self._order_by = tuple(order_by)

    def partitionBy(self, *cols) -> "WindowSpec":
        cols = cols[0] if len(cols) == 1 and isinstance(cols[0], (list, tuple)) else cols
        return WindowSpec(tuple(as_column(c) for c in cols), self._order_by)

    def orderBy(self, *cols) -> "WindowSpec":
        cols = cols[0] if len(cols) == 1 and isinstance(cols[0], (list, tuple)) else cols
        return WindowSpec(self._partition_by, tuple(as_column(c) for c in cols))


class Window:
    @staticmethod
    def partitionBy(*cols) -> WindowSpec:
        return WindowSpec().partitionBy(*cols)

    @staticmethod
    def orderBy(*cols) -> WindowSpec:
        return WindowSpec().orderBy(*cols)
