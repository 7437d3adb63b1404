"""Arrow-native DataFrame engine with a PySpark-shaped API.

It replaces the reference's JVM layer (``core/raydp-main``) "only as far as it feeds Arrow
record batches into the object store": partitions are Arrow tables, operators are
``pyarrow.compute`` kernels run on executor actors, and there is no JVM, py4j or row format.
"""
from raydp_b200.sql import functions, types  # noqa: F401
from raydp_b200.sql.column import Column  # noqa: F401
from raydp_b200.sql.dataframe import DataFrame, GroupedData, StorageLevel  # noqa: F401
from raydp_b200.sql.session import SparkConf, SparkContext, SparkSession  # noqa: F401
from raydp_b200.sql.types import Row  # noqa: F401
from raydp_b200.sql.window import Window  # noqa: F401
