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


def test_reference_raytrain_example_runs_unchanged(tmp_path):
    """``examples/raytrain_nyctaxi.py`` of the reference: first-generation ``ray.train.Trainer`` /
    ``TrainingCallback``, ``from ray import tune``, ``get_dataset_shard(...).to_torch(...)``."""
    import glob
    import shutil
    import pytest
    src = "/root/reference/examples"
    if not os.path.isdir(src):
        pytest.skip("reference examples not mounted")
    for f in glob.glob(os.path.join(src, "*.py")) + glob.glob(os.path.join(src, "*.csv")):
        shutil.copy(f, tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    proc = subprocess.run([sys.executable, "-m", "raydp_b200.compat", "raytrain_nyctaxi.py"], cwd=str(tmp_path),
                          env=env, capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-2000:]


def test_reference_tutorial_notebook_code_runs(tmp_path):
    """The code cells of ``tutorials/pytorch_example.ipynb`` (shell cells dropped, the dataset path pointed
    at the file in the reference tree): rollup, describe, ``avg`` over a text column with ``N/A`` entries,
    ``@udf`` cleaning, ``random_split``, ``TorchEstimator`` with per-column dtypes and torchmetrics."""
    import json
    import pytest
    nb = "/root/reference/tutorials/pytorch_example.ipynb"
    data = "/root/reference/tutorials/dataset/healthcare-dataset-stroke-data.csv"
    if not (os.path.exists(nb) and os.path.exists(data)):
        pytest.skip("reference tutorials not mounted")
    cells = json.load(open(nb))["cells"]
    code = "\n\n".join("".join(c["source"]) for c in cells if c["cell_type"] == "code")
    code = "\n".join(line for line in code.splitlines() if not line.startswith("!"))
    code = code.replace("/content/healthcare-dataset-stroke-data.csv", data)
    script = tmp_path / "pytorch_example.py"
    script.write_text(code)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    proc = subprocess.run([sys.executable, "-m", "raydp_b200.compat", str(script)], cwd=str(tmp_path), env=env,
                          capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-2000:]
    assert "|  null| 5110|" in proc.stdout                        # the rollup totals of the 5110-row dataset


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
            "'test_from_spark.py', 'test_spark_master_memory.py', 'test_mpi.py']))")
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    proc = subprocess.run([sys.executable, "-c", code], cwd=str(tmp_path), env=env, capture_output=True,
                          text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout[-3000:] + proc.stderr[-2000:]
    assert "15 passed" in proc.stdout        # test_mpi's 10 s limits also guard the start-up time of worker processes


_DOC_SNIPPETS = '''
import json
import ray
import raydp
ray.init(num_cpus=4, resources={"spark_master": 1, "spark_executor": 2})
# doc/spark_on_ray.md, spark-submit recipe: the node facts that go into ray.conf
node = ray.worker.global_worker.node
options = {"ray": {"run-mode": "CLUSTER", "node-ip": node.node_ip_address, "address": node.address,
                   "session-dir": node.get_session_dir_path()}}
json.dumps(options)
# doc/spark_on_ray.md: actor resources for master and executors, oversubscription, shuffle service, hive
spark = raydp.init_spark(app_name="RayDP Oversubscribe Example", num_executors=1, executor_cores=3,
                         executor_memory=1 * 1024 * 1024 * 1024,
                         configs={"spark.ray.raydp_spark_executor.actor.resource.cpu": 1,
                                  "spark.ray.raydp_spark_master.actor.resource.CPU": 0,
                                  "spark.ray.raydp_spark_master.actor.resource.spark_master": 1,
                                  "spark.ray.raydp_spark_executor.actor.resource.spark_executor": 1,
                                  "spark.shuffle.service.enabled": "true",
                                  "raydp.executor.extraClassPath": "/your/extra/jar/path:/another/path"},
                         enable_hive=True)
free = ray.available_resources()
assert free["CPU"] == 3.0 and free["spark_executor"] == 1.0 and "spark_master" not in free, free
assert spark.conf.get("spark.sql.warehouse.dir", None) is None
spark.sql("create database if not exists db")
spark.range(10).write.saveAsTable("db.xxx")
assert spark.sql("select * from db.xxx").count() == 10
# README.md: word count, Spark <-> dataset conversion, recoverable conversion
df = spark.createDataFrame([("look",), ("spark",), ("tutorial",), ("spark",), ("look",), ("python",)], ["word"])
counts = {r["word"]: r["count"] for r in df.groupBy("word").count().collect()}
assert counts == {"look": 2, "spark": 2, "tutorial": 1, "python": 1}, counts
ds1 = ray.data.from_spark(spark.range(0, 1000))
ds2 = ray.data.from_items([{"id": i} for i in range(1000)])
assert ds1.count() == 1000 and ds2.to_spark(spark).count() == 1000
assert raydp.spark.from_spark_recoverable(spark.range(100000)).count() == 100000
raydp.stop_spark()
ray.shutdown()
print("doc snippets OK")
'''


def test_reference_doc_snippets_run_through_compat(tmp_path):
    """The code blocks of the reference's README.md and doc/spark_on_ray.md (custom actor resources,
    oversubscription, shuffle service, hive tables, ``ray.worker.global_worker.node`` for ray.conf)."""
    script = tmp_path / "snippets.py"
    script.write_text(_DOC_SNIPPETS)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    proc = subprocess.run([sys.executable, "-m", "raydp_b200.compat", str(script)], cwd=str(tmp_path), env=env,
                          capture_output=True, text=True, timeout=420)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-2000:]
    assert "doc snippets OK" in proc.stdout


def test_reference_xgboost_ray_example_runs_unchanged(tmp_path):
    """``examples/xgboost_ray_nyctaxi.py`` of the reference talks to ``xgboost_ray`` directly (``RayDMatrix``,
    ``train``, ``RayParams``); with the alias in place it boosts on this framework's distributed booster."""
    import glob
    import shutil
    import pytest
    src = "/root/reference/examples"
    if not os.path.isdir(src):
        pytest.skip("reference examples not mounted")
    for f in glob.glob(os.path.join(src, "*.py")) + glob.glob(os.path.join(src, "*.csv")):
        shutil.copy(f, tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    proc = subprocess.run([sys.executable, "-m", "raydp_b200.compat", "xgboost_ray_nyctaxi.py"], cwd=str(tmp_path),
                          env=env, capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-2000:]
    assert "Final validation error:" in proc.stdout


def test_reference_dlrm_notebook_code_runs(tmp_path):
    """The code cells of ``examples/pytorch_dlrm.ipynb``: Criteo preprocessing on the DataFrame engine
    (``posexplode(array(...))``, window-assigned category ids, broadcast joins, log transforms), the notebook's own
    ``Dlrm`` module, ``TorchEstimator(num_workers=2)`` with 13 float + 26 long feature columns.  Only the data
    path (a generated Criteo-format TSV), the proxy ``del os.environ`` lines and the executor size (2 x 4 cores do
    not leave room for two training workers on an 8-core box) are adjusted."""
    import json
    import numpy as np
    import pytest
    nb = "/root/reference/examples/pytorch_dlrm.ipynb"
    if not os.path.exists(nb):
        pytest.skip("reference examples not mounted")
    cells = json.load(open(nb))["cells"]
    code = "\n\n".join("".join(c["source"]) for c in cells if c["cell_type"] == "code")
    code = "\n".join(line for line in code.splitlines()
                     if not line.startswith("!") and not line.startswith("del os.environ"))
    data = tmp_path / "criteo.tsv"
    assert "/Users/xianyang/datasets/criteo-small/train_1m.txt" in code and "executor_cores=4," in code
    code = code.replace("/Users/xianyang/datasets/criteo-small/train_1m.txt", str(data))
    code = code.replace("executor_cores=4,", "executor_cores=2,")
    rng = np.random.default_rng(0)
    with open(data, "w") as fh:
        for _ in range(2000):
            ints = ["" if rng.random() < 0.1 else str(int(rng.integers(0, 100))) for _ in range(13)]
            cats = ["" if rng.random() < 0.05 else format(int(rng.integers(0, 50 + 20 * j)), "08x") for j in range(26)]
            fh.write("\t".join([str(int(rng.random() < 0.25))] + ints + cats) + "\n")
    script = tmp_path / "pytorch_dlrm.py"
    script.write_text(code)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    proc = subprocess.run([sys.executable, "-m", "raydp_b200.compat", str(script)], cwd=str(tmp_path), env=env,
                          capture_output=True, text=True, timeout=900)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-3000:]


def test_reference_horovod_example_runs_unchanged(tmp_path):
    """``examples/horovod_nyctaxi.py`` of the reference: ``import horovod.torch as hvd`` (init / rank / size,
    ``broadcast_parameters``, ``broadcast_optimizer_state``, ``DistributedOptimizer(op=hvd.Average)``) and
    ``ray.train.horovod.HorovodTrainer``; the workers' progress lines reach the driver as with Ray."""
    import glob
    import shutil
    import pytest
    src = "/root/reference/examples"
    if not os.path.isdir(src):
        pytest.skip("reference examples not mounted")
    for f in glob.glob(os.path.join(src, "*.py")) + glob.glob(os.path.join(src, "*.csv")):
        shutil.copy(f, tmp_path)
    env = dict(os.environ, PYTHONPATH=ROOT + os.pathsep + os.environ.get("PYTHONPATH", ""))
    proc = subprocess.run([sys.executable, "-m", "raydp_b200.compat", "horovod_nyctaxi.py", "--epochs", "2"],
                          cwd=str(tmp_path), env=env, capture_output=True, text=True, timeout=600)
    assert proc.returncode == 0, proc.stdout[-2000:] + proc.stderr[-2000:]
    lines = [ln for ln in proc.stderr.splitlines() if "Train Epoch: 2" in ln]
    assert len(lines) == 2 and all(ln.startswith("(actor-") for ln in lines), proc.stderr[-1500:]
