"""Packaging for raydp_b200.

The reference's ``python/setup.py`` (``:28-146``) drives Maven, copies jars into the wheel and
generates gRPC stubs.  There is no JVM here: the native parts are two C-ABI shared libraries --
the shared-memory object store (g++) and the sm_100a kernels + peer-memory communicator (nvcc) --
which ``build_py`` compiles in place so they ship inside the wheel next to
``raydp_b200/_native/__init__.py``.  ``RAYDP_B200_SKIP_CUDA=1`` builds a CPU-only wheel
(DataFrame engine, runtime, store, CPU training) on machines without the CUDA toolkit.

    python setup.py build_py          # compile in place
    python -m pip install --no-build-isolation --no-deps .
"""
import os
import sys

from setuptools import find_packages, setup
from setuptools.command.build_py import build_py as _build_py

ROOT = os.path.dirname(os.path.abspath(__file__))


def _version() -> str:
    scope = {}
    with open(os.path.join(ROOT, "raydp_b200", "__init__.py")) as f:
        for line in f:
            if line.startswith("__version__"):
                exec(line, scope)
    return scope["__version__"]


class build_py(_build_py):
    """Compile the native libraries before the Python files are collected."""

    def run(self):
        sys.path.insert(0, ROOT)
        from raydp_b200 import _native
        _native.build_store(verbose=True)
        if os.environ.get("RAYDP_B200_SKIP_CUDA") != "1":
            if _native.nvcc_path() is None:
                raise RuntimeError("nvcc not found; set RAYDP_B200_SKIP_CUDA=1 for a CPU-only build")
            _native.build_kernels(verbose=True)
        super().run()


setup(
    name="raydp-b200",
    version=_version(),
    description="Blackwell-native Spark-to-GPU training framework with RayDP's public API",
    long_description=open(os.path.join(ROOT, "README.md")).read(),
    long_description_content_type="text/markdown",
    packages=find_packages(include=["raydp_b200", "raydp_b200.*", "raydp"]),
    package_data={"raydp_b200._native": ["*.so"]},
    data_files=[("include", ["include/raydp_b200.h", "include/raydp_b200_store.h"])],
    scripts=["bin/raydp-submit"],
    python_requires=">=3.10",
    install_requires=["numpy", "pyarrow", "pandas", "cloudpickle", "psutil", "torch"],
    extras_require={"mpi": ["grpcio", "protobuf"], "xgboost": ["xgboost"], "tf": ["tensorflow"]},
    cmdclass={"build_py": build_py},
    zip_safe=False,
)
