"""Script for ``bin/raydp-submit``: the session comes from the submit configuration.

    bin/raydp-submit --conf spark.executor.instances=2 --conf spark.executor.cores=1 \
        --conf spark.executor.memory=500M examples/raydp_submit_example.py
"""
from raydp_b200.sql import SparkSession

spark = SparkSession.builder.appName("submitted").getOrCreate()
df = spark.range(0, 1000, numPartitions=4)
print("count =", df.count(), "executors =", spark.conf.get("spark.executor.instances"))
