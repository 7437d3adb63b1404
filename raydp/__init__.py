"""Drop-in alias: ``import raydp`` resolves to :mod:`raydp_b200`.

Code written against the reference (``raydp.init_spark``, ``raydp.torch.TorchEstimator``,
``raydp.spark.spark_dataframe_to_ray_dataset``, ``raydp.utils.random_split``, ``raydp.mpi`` ...)
imports unchanged.  Sub-modules are aliased lazily through a meta-path finder so that
``import raydp.torch.estimator`` and ``raydp_b200.torch.estimator`` are the same module object.
"""
import importlib
import importlib.abc
import importlib.util
import sys

import raydp_b200 as _impl
from raydp_b200 import __version__, init_spark, stop_spark  # noqa: F401


class _AliasFinder(importlib.abc.MetaPathFinder, importlib.abc.Loader):
    _prefix = "raydp."

    def find_spec(self, fullname, path=None, target=None):
        if not fullname.startswith(self._prefix):
            return None
        real = "raydp_b200." + fullname[len(self._prefix):]
        try:
            if importlib.util.find_spec(real) is None:
                return None
        except (ImportError, ValueError):
            return None
        return importlib.util.spec_from_loader(fullname, self)

    def create_module(self, spec):
        real = "raydp_b200." + spec.name[len(self._prefix):]
        return importlib.import_module(real)

    def exec_module(self, module):
        return None


if not any(isinstance(f, _AliasFinder) for f in sys.meta_path):
    sys.meta_path.insert(0, _AliasFinder())

# Scripts written against the reference also ``import ray`` and ``from pyspark.sql...``: provide those
# names (pointing at this framework's runtime / DataFrame engine) unless the real packages exist.
from raydp_b200 import compat as _compat  # noqa: E402

_compat.install()

__all__ = ["init_spark", "stop_spark", "__version__"]


def __getattr__(name):
    try:
        return importlib.import_module("raydp." + name)
    except ImportError as exc:
        raise AttributeError(name) from exc
