"""Spark-on-runtime layer: cluster objects and DataFrame <-> Dataset exchange.

Parity: reference ``python/raydp/spark/__init__.py:18-36``.
"""
from raydp_b200.spark.dataset import (PartitionObjectsOwner, from_spark_recoverable,  # noqa: F401
                                      get_raydp_master_owner, ray_dataset_to_spark_dataframe,
                                      read_spark_parquet, release_spark_recoverable,
                                      spark_dataframe_to_ray_dataset)
from raydp_b200.spark.interfaces import SparkEstimatorInterface  # noqa: F401
from raydp_b200.spark.ray_cluster import SparkCluster  # noqa: F401

__all__ = ["SparkCluster", "SparkEstimatorInterface", "PartitionObjectsOwner",
           "get_raydp_master_owner", "spark_dataframe_to_ray_dataset", "from_spark_recoverable",
           "release_spark_recoverable", "ray_dataset_to_spark_dataframe", "read_spark_parquet"]
