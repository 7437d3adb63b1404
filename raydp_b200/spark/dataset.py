"""Spark DataFrame <-> Dataset exchange as Arrow blocks through the object store.

Parity: reference ``python/raydp/spark/dataset.py:47-320``.

* ``from_spark_recoverable`` (``:172-222``): repartition if asked, persist the partitions as
  Arrow in the executors' block managers and pin the lineage with the driver agent
  (``ObjectStoreWriter.prepareRecoverableRDD`` ``ObjectStoreWriter.scala:164-204``), then one
  fetch task per partition with ``max_retries=-1`` (``:54-84``) calling
  ``raydp-executor-<id>.get_rdd_partition`` (``RayDPExecutor.getRDDPartition``); fetch-task
  resources come from ``spark.ray.raydp_recoverable_fetch.task.resource.{CPU,memory}``
  (``:195-200``).  Recovery = actor restart + task retry + recache from lineage.
* ``spark_dataframe_to_ray_dataset`` (``:161-170``; ``owner`` ignored with a warning).
* ``release_spark_recoverable`` (``:224-237``).
* ``ray_dataset_to_spark_dataframe`` (``:293-313``): Arrow-typed datasets only; ``bytes``
  blocks (IPC streams) and ``pa.Table`` blocks both supported; refs are parked in the master
  actor (object holder) so they outlive the caller.
* ``PartitionObjectsOwner`` / ``get_raydp_master_owner`` (``:137-158``), ``read_spark_parquet``
  (``:316-320``).
"""
from __future__ import annotations

import logging
import uuid
import weakref
from dataclasses import dataclass
from typing import Any, Callable, Dict, List, Optional, Union

import pyarrow as pa

from raydp_b200 import data as rdata
from raydp_b200 import runtime
from raydp_b200.spark import configs as C
from raydp_b200.sql import types as T
from raydp_b200.sql.dataframe import DataFrame, StorageLevel
from raydp_b200.sql.plan import Plan, RefSource
from raydp_b200.sql.session import SparkSession

logger = logging.getLogger(__name__)

# DataFrame -> recoverable rdd id (weak keys: an unreferenced DataFrame cannot be released
# explicitly any more; its RDD goes away with the session)
_recoverable_rdd_ids: "weakref.WeakKeyDictionary[DataFrame, int]" = weakref.WeakKeyDictionary()


@runtime.remote(max_retries=-1)
def _fetch_arrow_table_from_executor(executor_actor_name: str, rdd_id: int, partition_id: int):
    """Fetch-task body: pull one cached Arrow partition out of an executor actor.

    The returned table becomes a block in the object store with this task as its lineage, so a
    lost block is rebuilt by re-running the fetch against the (restarted) executor.
    """
    executor = runtime.get_actor(executor_actor_name)
    table = runtime.get(executor.get_rdd_partition.remote(rdd_id, partition_id))
    return table.replace_schema_metadata()


@dataclass
class PartitionObjectsOwner:
    # name of the actor that should own the converted partitions
    actor_name: str
    # callable(actor_handle, refs) -> ObjectRef storing the refs in the actor's state
    set_reference_as_state: Callable[[Any, List[Any]], Any]


def get_raydp_master_owner(spark: Optional[SparkSession] = None) -> PartitionObjectsOwner:
    if spark is None:
        spark = SparkSession.getActiveSession()
    holder_name = spark.sparkContext.appName + C.RAYDP_SPARK_MASTER_SUFFIX

    def set_reference_as_state(master_actor, objects: List[Any]):
        return master_actor.add_objects.remote(str(uuid.uuid4()), objects)

    return PartitionObjectsOwner(holder_name, set_reference_as_state)


def spark_dataframe_to_ray_dataset(df: DataFrame, parallelism: Optional[int] = None,
                                   owner: Union[PartitionObjectsOwner, None] = None) -> rdata.Dataset:
    """Convert a Spark DataFrame to a Dataset (recoverable pipeline)."""
    if owner is not None:
        logger.warning("spark_dataframe_to_ray_dataset now uses recoverable conversion by default; "
                       "the 'owner' argument is ignored.")
    return from_spark_recoverable(df, parallelism=parallelism)


def from_spark_recoverable(df: DataFrame, storage_level: StorageLevel = StorageLevel.MEMORY_AND_DISK,
                           parallelism: Optional[int] = None) -> rdata.Dataset:
    """Recoverable Spark -> Dataset conversion that survives executor loss."""
    num_part = df.rdd.getNumPartitions()
    if parallelism is not None and parallelism != num_part:
        df = df.repartition(parallelism)
    df._materialise_cache()
    session = df.sparkSession
    backend = session._backend
    schema = df._arrow_schema()
    if not hasattr(backend, "master"):
        # local[n] session: no executors to lose, blocks are simply put into the store
        tables = backend.run(df._plan, "collect", {})
        return rdata.from_arrow([t.replace_schema_metadata() for t in tables])

    master = backend.master
    rdd_id = runtime.get(master.new_rdd_id.remote())
    runtime.get(master.pin_recoverable.remote(rdd_id, df._plan))
    actor_ids = backend.executor_actor_ids()
    n = df._plan.num_partitions
    preferred = {i: i % len(actor_ids) for i in range(n)}
    stats = backend.run(df._plan, "cache", {"rdd_id": rdd_id}, preferred=preferred)
    locations = [actor_ids[preferred[i]] for i in range(n)]

    conf = session.sparkContext.getConf()
    prefix = C.RECOVERABLE_FETCH_TASK_RESOURCE_PREFIX
    fetch_cpus = float(conf.get(prefix + ".CPU", "0"))
    fetch_memory = float(conf.get(prefix + ".memory", "0"))
    app = runtime.get(master.application_info.remote(backend.app_id))
    node_of = {info["actor"]: info["node_ip"] for info in app.get("executors", {}).values()}

    def submit(i):
        # run the fetch next to the executor that caches the partition (block locality)
        # (a soft preference: after a node loss the retry may run anywhere)
        task = _fetch_arrow_table_from_executor.options(
            num_cpus=fetch_cpus, memory=fetch_memory or None,
            _prefer_node_ip=node_of.get(locations[i]))
        return task.remote(C.EXECUTOR_NAME_PREFIX + str(locations[i]), rdd_id, i)
    refs = rdata._in_waves(submit, n)
    runtime.wait(refs, num_returns=len(refs), timeout=None)
    _recoverable_rdd_ids[df] = rdd_id
    # the master actor (object holder) owns the converted blocks: they die with
    # stop_spark(cleanup_data=True) and survive stop_spark(cleanup_data=False)
    runtime.get(master.add_objects.remote(f"rdd-{rdd_id}-{uuid.uuid4().hex[:8]}", list(refs)))
    metadata = [rdata.BlockMetadata(rows, nbytes, schema) for rows, nbytes in stats]
    return rdata.from_arrow_refs(refs, metadata)


def release_spark_recoverable(df: DataFrame) -> None:
    """Drop the cached partitions and the pinned lineage made by ``from_spark_recoverable``.
    Without this call they are released when the session stops."""
    rdd_id = _recoverable_rdd_ids.pop(df, None)
    if rdd_id is None:
        return
    backend = df.sparkSession._backend
    if hasattr(backend, "master"):
        runtime.get(backend.master.release_recoverable.remote(rdd_id))
        runtime.get([ex.drop_rdd.remote(rdd_id) for ex in backend.executors()])


def get_locations(blocks: List[Any]) -> List[Optional[str]]:
    """Node IP holding each block (locality hint; reference ``:285-291`` asks the core worker
    for owner addresses)."""
    return runtime.object_locations(blocks)


def _convert_by_rdd(spark: SparkSession, blocks: List[Any], locations, schema: pa.Schema) -> DataFrame:
    """``bytes`` blocks = Arrow IPC streams; decoded per partition on the executors."""
    return DataFrame(spark, Plan(RefSource([[b] for b in blocks], schema)))


def _convert_by_udf(spark: SparkSession, blocks: List[Any], locations, schema: pa.Schema) -> DataFrame:
    """``pa.Table`` blocks: park the refs in the master actor so they stay alive, then read them
    in place on the executors (no pandas round trip as in reference ``:239-269``)."""
    holder_name = spark.sparkContext.appName + C.RAYDP_SPARK_MASTER_SUFFIX
    try:
        holder = runtime.get_actor(holder_name)
        runtime.get(holder.add_objects.remote(str(uuid.uuid4()), list(blocks)))
    except ValueError:
        pass  # local[n] session has no holder actor; the caller keeps the refs alive
    return DataFrame(spark, Plan(RefSource([[b] for b in blocks], schema)))


def ray_dataset_to_spark_dataframe(spark: SparkSession, arrow_schema, blocks: List[Any],
                                   locations=None) -> DataFrame:
    locations = get_locations(blocks)
    if hasattr(arrow_schema, "base_schema"):
        arrow_schema = arrow_schema.base_schema
    if not isinstance(arrow_schema, pa.Schema):
        raise RuntimeError(f"Schema is {type(arrow_schema)}, required pyarrow.lib.Schema. \n"
                           f"to_spark does not support converting non-arrow ray datasets.")
    for field in arrow_schema:      # every field must map onto a Spark SQL type
        T.from_arrow_type(field.type)
    sample = runtime.get(blocks[0])
    if isinstance(sample, (bytes, bytearray, memoryview, pa.Buffer)):
        return _convert_by_rdd(spark, blocks, locations, arrow_schema)
    if isinstance(sample, pa.Table):
        return _convert_by_udf(spark, blocks, locations, arrow_schema)
    raise RuntimeError("ray.to_spark only supports arrow type blocks")


def read_spark_parquet(path: str) -> rdata.Dataset:
    """Read a directory written by ``df.write.parquet`` (skips ``_SUCCESS`` / crc files)."""
    return rdata.read_parquet(path, file_extensions=["parquet"])
