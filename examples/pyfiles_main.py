"""Driver script for the ``--py-files`` check of ``bin/raydp-submit``.

    bin/raydp-submit --py-files examples/helper_pyfile.py \
        --conf spark.executor.instances=1 --conf spark.executor.cores=1 \
        --conf spark.executor.memory=500M examples/pyfiles_main.py

Counterpart of the reference's ``examples/test_pyfiles_main.py``: the helper module must import
on the driver and inside functions that run on the executors (``mapPartitions``).
"""
import sys

from raydp_b200.sql import SparkSession

try:
    import helper_pyfile
except ImportError as exc:                                   # pragma: no cover - failure path
    print("FAILED: helper_pyfile is not importable on the driver:", exc)
    sys.exit(1)


def per_partition(rows):
    import helper_pyfile as shipped                          # resolved on the executor
    yield shipped.describe(shipped.scale(r[0]) for r in rows)


spark = SparkSession.builder.appName("py-files check").getOrCreate()
print("driver-side call:", helper_pyfile.scale(14))
data = spark.sparkContext.parallelize([(i,) for i in range(1, 11)], 2)
parts = data.mapPartitions(per_partition).collect()
total = sum(p["sum"] for p in parts)
print("executor-side partitions:", parts)
assert helper_pyfile.scale(14) == 42 and total == 3 * 55, (parts, total)
print("py-files OK")
spark.stop()
