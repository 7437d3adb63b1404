"""Session lifecycle: ``init_spark`` / ``stop_spark``.

Parity: reference ``python/raydp/context.py:33-257`` -- one process-global context guarded by
an RLock (``:178-179``); ``init_spark`` auto-starts the runtime (``:215-217``), raises
``"The spark environment has inited."`` on double init (``:240``) and tears down on failure
(``:234-238``); placement groups are built from the strategy with one
``{"CPU": cores, "memory": bytes}`` bundle per executor (``:119-140``) and recorded in
``spark.ray.placement_group`` / ``spark.ray.bundle_indexes``; ``stop_spark(cleanup_data)``
keeps the master actor -- and the data it holds -- alive when ``cleanup_data=False``
(``:243-254``); ``atexit`` registration (``:257``).
"""
from __future__ import annotations

import atexit
from contextlib import ContextDecorator
from threading import RLock
from typing import Dict, List, Optional, Union

from raydp_b200 import runtime
from raydp_b200.spark.ray_cluster import SparkCluster
from raydp_b200.sql.session import SparkSession, reset_builder_singletons
from raydp_b200.utils import parse_memory_size


class _SparkContext(ContextDecorator):
    _PLACEMENT_GROUP_CONF = "spark.ray.placement_group"
    _BUNDLE_INDEXES_CONF = "spark.ray.bundle_indexes"

    def __init__(self, app_name: str, num_executors: int, executor_cores: int,
                 executor_memory: Union[str, int], enable_hive: bool, fault_tolerant_mode: bool,
                 placement_group_strategy: Optional[str], placement_group,
                 placement_group_bundle_indexes: Optional[List[int]],
                 configs: Optional[Dict[str, str]] = None):
        self._app_name = app_name
        self._num_executors = num_executors
        self._executor_cores = executor_cores
        self._executor_memory = executor_memory
        self._enable_hive = enable_hive
        self._fault_tolerant_mode = fault_tolerant_mode
        self._placement_group_strategy = placement_group_strategy
        self._placement_group = placement_group
        self._placement_group_bundle_indexes = placement_group_bundle_indexes
        self._configs: Optional[Dict[str, str]] = {} if configs is None else dict(configs)
        self._spark_cluster: Optional[SparkCluster] = None
        self._spark_session: Optional[SparkSession] = None

    def _get_or_create_spark_cluster(self) -> SparkCluster:
        if self._spark_cluster is None:
            self._spark_cluster = SparkCluster(self._app_name, self._num_executors,
                                               self._executor_cores, self._executor_memory,
                                               self._enable_hive, self._configs)
        return self._spark_cluster

    def _prepare_placement_group(self) -> None:
        if self._placement_group_strategy is not None:
            memory = self._executor_memory
            if isinstance(memory, str):
                memory = parse_memory_size(memory)
            bundles = [{"CPU": self._executor_cores, "memory": memory}
                       for _ in range(self._num_executors)]
            pg = runtime.placement_group(bundles, strategy=self._placement_group_strategy)
            runtime.get(pg.ready())
            self._placement_group = pg
            self._placement_group_bundle_indexes = None
        if self._placement_group is not None:
            self._configs[self._PLACEMENT_GROUP_CONF] = self._placement_group.id.hex()
            idxs = (list(range(self._placement_group.bundle_count))
                    if self._placement_group_bundle_indexes is None
                    else self._placement_group_bundle_indexes)
            self._configs[self._BUNDLE_INDEXES_CONF] = ",".join(map(str, idxs))

    def get_or_create_session(self) -> SparkSession:
        if self._spark_session is not None:
            return self._spark_session
        self._prepare_placement_group()
        cluster = self._get_or_create_spark_cluster()
        self._spark_session = cluster.get_spark_session()
        if self._fault_tolerant_mode:
            cluster.connect_spark_driver_to_ray()
        return self._spark_session

    def stop(self, cleanup_data: bool = True) -> None:
        if self._spark_session is not None:
            self._spark_session.stop()
            self._spark_session = None
            reset_builder_singletons()
        if self._spark_cluster is not None:
            self._spark_cluster.stop(cleanup_data)
            if cleanup_data:
                self._spark_cluster = None
        if self._placement_group_strategy is not None:
            if self._placement_group is not None:
                runtime.remove_placement_group(self._placement_group)
                self._placement_group = None
            self._placement_group_strategy = None
        if self._configs is not None:
            self._configs = None

    def __enter__(self):
        self.get_or_create_session()

    def __exit__(self, exc_type, exc_val, exc_tb):
        self.stop()


_spark_context_lock = RLock()
_global_spark_context: Optional[_SparkContext] = None


def init_spark(app_name: str, num_executors: int, executor_cores: int,
               executor_memory: Union[str, int], enable_hive: bool = False,
               fault_tolerant_mode: bool = True, placement_group_strategy: Optional[str] = None,
               placement_group=None, placement_group_bundle_indexes: Optional[List[int]] = None,
               configs: Optional[Dict[str, str]] = None) -> SparkSession:
    """Start the executors and return the Spark session.

    ``executor_memory`` accepts ``"500M"``-style strings or bytes.  ``fault_tolerant_mode``
    (default on) makes DataFrame -> Dataset conversion recoverable.  Either give a
    ``placement_group_strategy`` (PACK / SPREAD / STRICT_PACK / STRICT_SPREAD; a group is made
    and later removed for you) or an existing ``placement_group`` plus optional bundle indexes.
    """
    if not runtime.is_initialized():
        runtime.init()
    with _spark_context_lock:
        global _global_spark_context
        if _global_spark_context is None:
            try:
                _global_spark_context = _SparkContext(
                    app_name, num_executors, executor_cores, executor_memory, enable_hive,
                    fault_tolerant_mode, placement_group_strategy, placement_group,
                    placement_group_bundle_indexes, configs)
                return _global_spark_context.get_or_create_session()
            except BaseException:
                if _global_spark_context is not None:
                    try:
                        _global_spark_context.stop()
                    except Exception:  # noqa: BLE001
                        pass
                _global_spark_context = None
                raise
        else:
            raise Exception("The spark environment has inited.")


def stop_spark(cleanup_data: bool = True) -> None:
    """Stop the Spark session and executors.  With ``cleanup_data=False`` the master actor and
    the converted datasets it holds stay alive until the next ``stop_spark(True)``."""
    with _spark_context_lock:
        global _global_spark_context
        if _global_spark_context is not None:
            try:
                if runtime.is_initialized():
                    _global_spark_context.stop(cleanup_data)
            finally:
                if cleanup_data or not runtime.is_initialized():
                    _global_spark_context = None


atexit.register(stop_spark)
