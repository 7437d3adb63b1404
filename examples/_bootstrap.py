"""Make the examples runnable from a source checkout (``python examples/<name>.py``).

When the package is not pip-installed, the repository root is put on ``sys.path`` so that
``import raydp`` / ``import raydp_b200`` resolve; workers inherit the driver's path.
"""
import os
import sys

_ROOT = os.path.dirname(os.path.dirname(os.path.realpath(__file__)))
try:
    import raydp  # noqa: F401
except ModuleNotFoundError:
    sys.path.insert(0, _ROOT)
    import raydp  # noqa: F401
