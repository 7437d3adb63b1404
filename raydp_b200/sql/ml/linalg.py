"""``pyspark.ml.linalg``: dense and sparse vectors as thin numpy wrappers.  In DataFrames a vector column
is ``array<double>``; these classes are what ``Vectors.dense(...)`` literals and collected rows use."""
from __future__ import annotations

from typing import Iterable, Sequence, Union

import numpy as np


class DenseVector:
    def __init__(self, values: Iterable[float]):
        self.array = np.asarray(list(values) if not isinstance(values, np.ndarray) else values, dtype=np.float64)

    @property
    def values(self) -> np.ndarray:
        return self.array

    def toArray(self) -> np.ndarray:
        return self.array

    def tolist(self):
        return self.array.tolist()

    def dot(self, other) -> float:
        return float(np.dot(self.array, _as_array(other)))

    def norm(self, p: float = 2) -> float:
        return float(np.linalg.norm(self.array, p))

    def squared_distance(self, other) -> float:
        d = self.array - _as_array(other)
        return float(np.dot(d, d))

    def __len__(self) -> int:
        return len(self.array)

    def __getitem__(self, i):
        return self.array[i]

    def __iter__(self):
        return iter(self.array.tolist())

    def __eq__(self, other):
        return np.array_equal(self.array, _as_array(other))

    def __repr__(self):
        return f"DenseVector({self.array.tolist()})"


class SparseVector:
    def __init__(self, size: int, *args):
        self.size = int(size)
        if len(args) == 1:
            pairs = sorted(args[0].items()) if isinstance(args[0], dict) else sorted(args[0])
            self.indices = np.array([p[0] for p in pairs], dtype=np.int32)
            self.values = np.array([p[1] for p in pairs], dtype=np.float64)
        else:
            self.indices = np.asarray(args[0], dtype=np.int32)
            self.values = np.asarray(args[1], dtype=np.float64)

    def toArray(self) -> np.ndarray:
        out = np.zeros(self.size)
        out[self.indices] = self.values
        return out

    def tolist(self):
        return self.toArray().tolist()

    def __len__(self) -> int:
        return self.size

    def __eq__(self, other):
        return np.array_equal(self.toArray(), _as_array(other))

    def __repr__(self):
        return f"SparseVector({self.size}, {dict(zip(self.indices.tolist(), self.values.tolist()))})"


def _as_array(v) -> np.ndarray:
    if isinstance(v, (DenseVector, SparseVector)):
        return v.toArray()
    return np.asarray(v, dtype=np.float64)


class Vectors:
    @staticmethod
    def dense(*values) -> DenseVector:
        if len(values) == 1 and isinstance(values[0], (list, tuple, np.ndarray)):
            values = values[0]
        return DenseVector(values)

    @staticmethod
    def sparse(size: int, *args) -> SparseVector:
        return SparseVector(size, *args)

    @staticmethod
    def zeros(size: int) -> DenseVector:
        return DenseVector(np.zeros(size))

    @staticmethod
    def norm(vector, p: float) -> float:
        return float(np.linalg.norm(_as_array(vector), p))

    @staticmethod
    def squared_distance(a, b) -> float:
        d = _as_array(a) - _as_array(b)
        return float(np.dot(d, d))


Vector = Union[DenseVector, SparseVector, Sequence[float]]
