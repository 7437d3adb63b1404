"""``ParallelIteratorWorkerWithLen`` under its reference name.

Parity: reference ``python/raydp/spark/parallel_iterator_worker.py:23-29`` -- a shim over Ray's
removed ``ParallelIteratorWorker`` that adds ``__len__`` (dead code upstream).
Kept as a small self-contained actor-friendly iterator holder so imports of the name keep working.
"""
from __future__ import annotations

from typing import Any, Callable, Iterable, Iterator, List, Union


class ParallelIteratorWorkerWithLen:
    def __init__(self, item_generator: Union[Iterable[Any], Callable[[], Iterable[Any]]], repeat: bool,
                 num_records: int):
        self._source = item_generator
        self._repeat = repeat
        self.num_records = num_records
        self._transforms: List[Callable[[Iterator[Any]], Iterator[Any]]] = []
        self._live: Iterator[Any] = iter(())

    def _make(self) -> Iterator[Any]:
        base = self._source() if callable(self._source) else self._source
        it: Iterator[Any] = iter(base)
        for fn in self._transforms:
            it = fn(it)
        return it

    def par_iter_init(self, transforms: List[Callable[[Iterator[Any]], Iterator[Any]]]) -> None:
        self._transforms = list(transforms or [])
        self._live = self._make()

    def par_iter_next(self) -> Any:
        try:
            return next(self._live)
        except StopIteration:
            if not self._repeat:
                raise
            self._live = self._make()
            return next(self._live)

    def __len__(self) -> int:
        return self.num_records
