"""Estimator ABC.

Parity: reference ``python/raydp/estimator.py:23-43`` (``fit(train_ds, evaluate_ds=None)``,
``get_model()``).
"""
from abc import ABC, abstractmethod
from typing import Any, NoReturn, Optional


class EstimatorInterface(ABC):
    """scikit-learn-like distributed estimator."""

    @abstractmethod
    def fit(self, train_ds, evaluate_ds: Optional[Any] = None) -> NoReturn:
        """Train (and optionally evaluate) on :class:`raydp_b200.data.Dataset` objects."""

    @abstractmethod
    def get_model(self) -> Any:
        """Return the trained model."""
