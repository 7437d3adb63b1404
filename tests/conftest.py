"""Shared fixtures.

Mirrors the reference's fixture set (``python/raydp/tests/conftest.py:66-140``): a local Spark
session, a runtime ("ray cluster") and small Spark-on-runtime sessions; GPU tests carry the
``gpu`` marker and are deselected on CPU boxes with ``-m "not gpu"``.
"""
import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
os.environ.setdefault("RAYDP_B200_SPMD", "0")   # tests drive workers as actors, not torchrun


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on a B200 with -m gpu)")
    config.addinivalue_line("markers", "multigpu: needs >= 2 CUDA devices")


def pytest_collection_modifyitems(config, items):
    # the no-allocation entry point test must run after everything else
    items.sort(key=lambda it: 1 if "test_zz_" in it.nodeid else 0)


@pytest.fixture(scope="session")
def spark_session(tmp_path_factory):
    from raydp_b200.sql import SparkSession
    warehouse = str(tmp_path_factory.mktemp("warehouse"))       # tables of the tests stay out of the repo
    spark = SparkSession.builder.master("local[2]").appName("RayDP test") \
        .config("spark.sql.warehouse.dir", warehouse).getOrCreate()
    yield spark
    spark.stop()


@pytest.fixture(scope="function")
def ray_cluster():
    from raydp_b200 import runtime
    runtime.shutdown()
    runtime.init(address="local", num_cpus=6, include_dashboard=False)
    yield
    import raydp_b200
    raydp_b200.stop_spark()
    runtime.shutdown()


@pytest.fixture(scope="function")
def spark_on_ray_small(ray_cluster):
    import raydp_b200
    spark = raydp_b200.init_spark("test", 1, 1, "500M")
    yield spark
    raydp_b200.stop_spark()


@pytest.fixture(scope="function")
def spark_on_ray_2_executors(ray_cluster):
    import raydp_b200
    spark = raydp_b200.init_spark("test", 2, 1, "500M")
    yield spark
    raydp_b200.stop_spark()


def cuda_device_count() -> int:
    try:
        import torch
        return torch.cuda.device_count() if torch.cuda.is_available() else 0
    except Exception:  # noqa: BLE001
        return 0
