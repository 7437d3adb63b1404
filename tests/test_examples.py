"""The example scripts run as shipped (reference CI runs its examples the same way,
``.github/workflows/raydp.yml:112-125``)."""
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _run(args, timeout=420):
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    proc = subprocess.run(args, cwd=ROOT, env=env, capture_output=True, text=True, timeout=timeout)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-2000:]
    return proc.stdout


def test_data_process_example_builds_20_features(tmp_path):
    out = _run([sys.executable, "examples/data_process.py"])
    assert "features: 20" in out and "pickup_distance_jfk" in out


def test_pytorch_nyctaxi_example_trains():
    out = _run([sys.executable, "examples/pytorch_nyctaxi.py", "--epochs", "2", "--workers", "2"])
    assert '"example": "pytorch_nyctaxi"' in out and "NYC_Model parameters: 48513" in out


def test_submit_py_files_example():
    out = _run([os.path.join(ROOT, "bin", "raydp-submit"), "--py-files",
                "examples/helper_pyfile.py", "--conf", "spark.executor.instances=1",
                "--conf", "spark.executor.cores=1", "--conf", "spark.executor.memory=500M",
                "examples/pyfiles_main.py"])
    assert "py-files OK" in out


def test_reference_example_runs_unchanged_through_compat_runner():
    """A script written against the reference (``import ray``, ``from pyspark.sql.functions import ...``,
    ``raydp.init_spark`` with JVM-only options, ``ray.data.from_spark``) runs as is under
    ``python -m raydp_b200.compat``."""
    import pytest
    script = "/root/reference/examples/data_process.py"
    if not os.path.exists(script):
        pytest.skip("reference examples not mounted")
    proc_out = _run([sys.executable, "-m", "raydp_b200.compat", script])
    assert "Traceback" not in proc_out


def test_reference_unit_tests_pass_unchanged(tmp_path):
    """The reference's own pytest files (copied, not edited) run against this framework through the
    ``ray`` / ``pyspark`` aliases.  The quick ones run here; test_spark_cluster / test_torch / test_mpi /
    test_xgboost / test_read_parquet / test_torch_sequential pass the same way (see docs/COVERAGE.md)."""
    import glob
    import shutil
    import pytest
    src = "/root/reference/python/raydp/tests"
    if not os.path.isdir(src):
        pytest.skip("reference tests not mounted")
    for f in glob.glob(os.path.join(src, "*.py")):
        if os.path.basename(f) != "__init__.py":
            shutil.copy(f, tmp_path)
    code = ("from raydp_b200 import compat; compat.install(); import raydp, pytest, sys; "
            "sys.exit(pytest.main(['-x', '-q', '-p', 'no:cacheprovider', 'test_spark_utils.py', "
            "'test_from_spark.py', 'test_spark_master_memory.py']))")
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    proc = subprocess.run([sys.executable, "-c", code], cwd=str(tmp_path), env=env, capture_output=True,
                          text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-2000:]
    assert "7 passed" in proc.stdout
