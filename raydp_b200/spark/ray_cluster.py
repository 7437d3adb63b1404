"""Driver-side cluster object: creates the master actor and the Spark session.

Parity: reference ``python/raydp/spark/ray_cluster.py:34-210`` -- ``_prepare_spark_configs``
(``:115-189``: executor instances/cores/memory, driver host/bind address, dynamic-allocation
warning), ``_get_master_resources`` (``:80-98``: ``spark.ray.raydp_spark_master.actor.resource.*``
and the deprecated ``...master.resource.*``), ``_set_up_master`` (``:54-78``: named actor
``<app_name>_SPARK_MASTER``), ``get_spark_session`` (``:191-201``), ``stop(cleanup_data)``.
The classpath / javaagent / log4j plumbing has no meaning without a JVM and is dropped.
"""
from __future__ import annotations

import logging
from typing import Any, Dict, Optional, Union

from raydp_b200 import runtime
from raydp_b200.services import Cluster
from raydp_b200.spark import configs as C
from raydp_b200.spark.ray_cluster_master import RAYDP_SPARK_MASTER_SUFFIX, RayDPSparkMaster
from raydp_b200.utils import parse_memory_size

logger = logging.getLogger(__name__)


class SparkCluster(Cluster):
    def __init__(self, app_name: str, num_executors: int, executor_cores: int,
                 executor_memory: Union[str, int], enable_hive: bool,
                 configs: Optional[Dict[str, str]]):
        super().__init__(None)
        self._app_name = app_name
        self._spark_master = None
        self._num_executors = num_executors
        self._executor_cores = executor_cores
        self._executor_memory = executor_memory
        self._enable_hive = enable_hive
        self._configs: Dict[str, str] = {} if configs is None else dict(configs)
        self._prepare_spark_configs()
        self._set_up_master(resources=self._get_master_resources(self._configs), kwargs=None)
        self._spark_session = None

    # ------------------------------------------------------------------ master
    def _set_up_master(self, resources: Dict[str, float], kwargs: Optional[Dict[Any, Any]]):
        name = self._app_name + RAYDP_SPARK_MASTER_SUFFIX
        opts: Dict[str, Any] = {"name": name, "max_concurrency": 16}
        if resources:
            res = dict(resources)
            num_cpus = res.pop("CPU", None)
            memory = res.pop("memory", None)
            if num_cpus is not None:
                opts["num_cpus"] = num_cpus
            if memory is not None:
                opts["memory"] = memory
            opts["resources"] = res
        self._spark_master_handle = runtime.remote(RayDPSparkMaster).options(**opts).remote(self._configs)
        runtime.get(self._spark_master_handle.start_up.remote(resources))

    def _get_master_resources(self, configs: Dict[str, str]) -> Dict[str, float]:
        resources: Dict[str, float] = {}
        deprecated = C.resources_from_prefix(configs, C.SPARK_MASTER_RESOURCE_PREFIX_DEPRECATED)
        if deprecated:
            logger.warning("'%s.*' is deprecated, use '%s.*'",
                           C.SPARK_MASTER_RESOURCE_PREFIX_DEPRECATED,
                           C.SPARK_MASTER_ACTOR_RESOURCE_PREFIX)
            resources.update(deprecated)
        resources.update(C.resources_from_prefix(configs, C.SPARK_MASTER_ACTOR_RESOURCE_PREFIX))
        if "cpu" in resources:
            resources["CPU"] = resources.pop("cpu")
        return resources

    def _set_up_worker(self, resources: Dict[str, float], kwargs: Dict[str, str]):
        raise Exception("Unsupported operation")

    def get_cluster_url(self) -> str:
        return runtime.get(self._spark_master_handle.get_master_url.remote())

    def connect_spark_driver_to_ray(self) -> None:
        """Reference ``:106-113`` boots Ray inside the driver JVM for the recoverable path.
        The driver *is* this Python process and is already connected; nothing to do."""
        return None

    # ------------------------------------------------------------------ configs
    @staticmethod
    def _load_default_properties() -> Dict[str, str]:
        """``spark-defaults.conf`` from ``$SPARK_CONF_DIR`` or ``$SPARK_HOME/conf`` (reference: loaded by
        the app master, ``RayAppMaster.scala:61``; precedence user dict > file, ``doc/spark_on_ray.md:92``)."""
        import os
        for base in (os.environ.get("SPARK_CONF_DIR"),
                     os.path.join(os.environ["SPARK_HOME"], "conf") if os.environ.get("SPARK_HOME") else None):
            path = os.path.join(base, "spark-defaults.conf") if base else None
            if path and os.path.isfile(path):
                out: Dict[str, str] = {}
                with open(path) as f:
                    for line in f:
                        line = line.strip()
                        if not line or line.startswith("#"):
                            continue
                        parts = line.replace("=", " ", 1).split(None, 1)
                        if len(parts) == 2 and parts[0].startswith("spark."):
                            out[parts[0]] = parts[1].strip()
                return out
        return {}

    def _prepare_spark_configs(self) -> None:
        for key, value in self._load_default_properties().items():
            self._configs.setdefault(key, value)          # the user's dict wins
        cfg = self._configs
        cfg.setdefault("spark.executor.instances", str(self._num_executors))
        cfg["spark.executor.instances"] = str(self._num_executors)
        cfg["spark.executor.cores"] = str(self._executor_cores)
        memory = self._executor_memory
        if isinstance(memory, str):
            memory = parse_memory_size(memory)
        cfg["spark.executor.memory"] = str(int(memory))
        node_ip = runtime.get_node_ip_address()
        cfg.setdefault("spark.driver.host", node_ip)
        cfg.setdefault("spark.driver.bindAddress", node_ip)
        if self._enable_hive:
            cfg["spark.sql.catalogImplementation"] = "hive"
        dyn = str(cfg.get(C.DYN_ALLOC_ENABLED, "false")).lower() == "true"
        if dyn and C.DYN_ALLOC_MAX_EXECUTORS not in cfg:
            logger.warning("spark.dynamicAllocation.enabled is set without "
                           "spark.dynamicAllocation.maxExecutors: executors are only capped by "
                           "cluster resources; consider setting it")

    def get_spark_session(self):
        if self._spark_session is not None:
            return self._spark_session
        from raydp_b200.sql.session import SparkSession
        builder = SparkSession.builder
        for k, v in self._configs.items():
            builder = builder.config(k, v)
        if self._enable_hive:
            builder = builder.enableHiveSupport()
        self._spark_session = builder.appName(self._app_name).master(
            self.get_cluster_url()).getOrCreate()
        return self._spark_session

    def stop(self, cleanup_data: bool = True) -> None:
        if self._spark_session is not None:
            self._spark_session.stop()
            self._spark_session = None
        if self._spark_master_handle is not None:
            try:
                runtime.get(self._spark_master_handle.stop.remote(cleanup_data), timeout=60)
            except runtime.ActorDiedError:
                pass
            if cleanup_data:
                self._spark_master_handle = None
