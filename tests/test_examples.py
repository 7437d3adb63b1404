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
