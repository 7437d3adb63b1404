"""raydp_b200: a Blackwell-native Spark-to-GPU training framework with RayDP's public API.

Parity: reference ``python/raydp/__init__.py:18-22`` re-exports ``init_spark`` / ``stop_spark``
and ``__version__``.  ``import raydp`` (the thin alias package at the repo root) gives the
same names for code written against the reference.
"""
from raydp_b200.context import init_spark, stop_spark  # noqa: F401

__version__ = "1.7.0.dev0+b200"

__all__ = ["init_spark", "stop_spark", "__version__"]
