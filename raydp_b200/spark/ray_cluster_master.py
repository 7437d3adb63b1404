"""The ``<app_name>_SPARK_MASTER`` actor: app master + object holder + driver agent.

Parity (one Python actor replaces four reference components, because there is no JVM):

* ``RayDPSparkMaster`` (reference ``python/raydp/spark/ray_cluster_master.py:51-247``):
  ``start_up``, ``get_master_url``, ``get_ray_address``, ``get_actor_id``, the **object holder**
  ``add_objects(ts, objs)`` / ``get_object(ts, idx)`` (``:222-226``) that keeps converted data
  alive after Spark stops, and ``stop(cleanup_data)`` which only exits the actor when data is
  to be cleaned (``:236-247``).
* ``RayAppMaster`` (reference ``RayAppMaster.scala:42-371``): ``RegisterApplication`` ->
  schedule ``numExecutors`` executor actors (``:133-143,251-303``), ``RegisterExecutor`` /
  ``ExecutorStarted`` (``:151-171``), ``RequestExecutors`` / ``KillExecutors`` for dynamic
  allocation capped by ``spark.dynamicAllocation.maxExecutors`` (``:173-190,275-288``),
  ``RequestAddPendingRestartedExecutor`` with the new-id -> old-actor map (``:192-209``),
  round-robin over placement-group bundle indexes (``:341-349``).
* ``ApplicationInfo`` bookkeeping (reference ``ApplicationInfo.scala:33-184``).
* ``RayDPDriverAgent`` + ``ObjectStoreWriter.recoverableRDDs`` (reference
  ``RayDPDriverAgent.scala:25-78``, ``ObjectStoreWriter.scala:164-216``): the lineage of every
  recoverable RDD is pinned here so an executor that lost a block can recompute it.
"""
from __future__ import annotations

import logging
import os
import threading
import time
from typing import Any, Dict, List, Optional

from raydp_b200.exceptions import RayDPException
from raydp_b200.spark import configs as C

logger = logging.getLogger(__name__)

RAYDP_SPARK_MASTER_SUFFIX = C.RAYDP_SPARK_MASTER_SUFFIX


class ExecutorDesc:
    def __init__(self, executor_id: int, actor_executor_id: int):
        self.executor_id = executor_id
        self.actor_executor_id = actor_executor_id
        self.registered = False
        self.started = False
        self.node_ip: Optional[str] = None
        self.pid: Optional[int] = None
        self.handle = None


class ApplicationInfo:
    """Executor bookkeeping of one registered application."""

    def __init__(self, app_id: str, desc: Dict[str, Any]):
        self.id = app_id
        self.desc = desc
        self.start_time = time.time()
        self.executors: Dict[int, ExecutorDesc] = {}           # by Spark executor id
        self.by_actor: Dict[int, ExecutorDesc] = {}            # by actor id (stable name suffix)
        self.restarted_executors: Dict[int, int] = {}          # new executor id -> old id
        self.next_executor_id = 0
        self.state = "RUNNING"
        self.target_total = int(desc.get("num_executors", 0))

    def get_next_executor_id(self) -> int:
        eid = self.next_executor_id
        self.next_executor_id += 1
        return eid

    def add_pending(self, actor_executor_id: int) -> ExecutorDesc:
        ex = ExecutorDesc(actor_executor_id, actor_executor_id)
        self.executors[ex.executor_id] = ex
        self.by_actor[actor_executor_id] = ex
        return ex

    def register_executor(self, executor_id: int, node_ip: str, pid: int) -> bool:
        ex = self.executors.get(executor_id)
        if ex is None or ex.registered and ex.pid == pid:
            return ex is not None
        ex.registered, ex.node_ip, ex.pid = True, node_ip, pid
        return True

    def mark_started(self, executor_id: int) -> None:
        ex = self.executors.get(executor_id)
        if ex is not None:
            ex.started = True

    def remaining_unregistered(self) -> int:
        return sum(1 for e in self.executors.values() if not e.registered)

    def current_executors(self) -> int:
        return len(self.by_actor)

    def kill(self, actor_executor_id: int) -> Optional[ExecutorDesc]:
        ex = self.by_actor.pop(actor_executor_id, None)
        if ex is not None:
            self.executors.pop(ex.executor_id, None)
        return ex


class RayDPSparkMaster:
    """Actor body.  Created by :class:`raydp_b200.spark.ray_cluster.SparkCluster` under the
    name ``<app_name>_SPARK_MASTER`` with the resources from
    ``spark.ray.raydp_spark_master.actor.resource.*``."""

    def __init__(self, configs: Dict[str, str]):
        from raydp_b200 import runtime
        self._configs = dict(configs)
        self._lock = threading.RLock()
        self._apps: Dict[str, ApplicationInfo] = {}
        self._objects: Dict[str, List[Any]] = {}
        self._recoverable: Dict[int, Any] = {}       # rdd_id -> Plan (lineage), pinned
        self._next_rdd_id = 0
        self._name = runtime.get_runtime_context().get_actor_name()
        self._started = False
        self._app_seq = 0
        self._shuffle_services: Dict[str, Any] = {}   # node ip -> service actor handle

    # ------------------------------------------------------------------ lifecycle
    def start_up(self, resources: Optional[Dict[str, float]] = None) -> str:
        ignored = [k for k in self._configs if k in C.JVM_ONLY_KEYS]
        if ignored:
            logger.debug("JVM-only configuration keys accepted and ignored: %s", ignored)
        self._started = True
        return self.get_master_url()

    def get_master_url(self) -> str:
        return f"ray://{self._name}"

    def get_host(self) -> str:
        from raydp_b200 import runtime
        return runtime.get_node_ip_address()

    def get_ray_address(self) -> str:
        from raydp_b200 import runtime
        return runtime.session_info().get("address", "")

    def get_actor_id(self) -> Optional[str]:
        from raydp_b200 import runtime
        return runtime.get_runtime_context().get_actor_id()

    def get_spark_home(self) -> str:
        return os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

    # ------------------------------------------------------------------ object holder
    def add_objects(self, timestamp: str, objects: List[Any]) -> int:
        with self._lock:
            self._objects[timestamp] = list(objects)
        return len(objects)

    def get_object(self, timestamp: str, idx: int):
        # returning the ref's *value* keeps parity with the reference, whose holder returns
        # the ObjectRef and lets Ray resolve it
        from raydp_b200 import runtime
        return runtime.get(self._objects[timestamp][idx])

    def get_object_ref(self, timestamp: str, idx: int):
        return [self._objects[timestamp][idx]]

    def held_object_count(self) -> int:
        with self._lock:
            return sum(len(v) for v in self._objects.values())

    # ------------------------------------------------------------------ app master
    def register_application(self, desc: Dict[str, Any]) -> str:
        with self._lock:
            self._app_seq += 1
            app_id = f"app-{time.strftime('%Y%m%d%H%M%S')}-{self._app_seq:04d}"
            app = ApplicationInfo(app_id, desc)
            self._apps[app_id] = app
            for _ in range(app.target_total):
                self._request_new_executor(app)
        return app_id

    def _max_executors(self, app: ApplicationInfo) -> int:
        conf = app.desc.get("conf", {})
        if str(conf.get(C.DYN_ALLOC_ENABLED, "false")).lower() == "true":
            return int(conf.get(C.DYN_ALLOC_MAX_EXECUTORS, str(2 ** 31 - 1)))
        return int(conf.get("spark.executor.instances", app.desc.get("num_executors", 0)))

    def _request_new_executor(self, app: ApplicationInfo, force: bool = False) -> bool:
        from raydp_b200 import runtime
        from raydp_b200.sql.executor import SparkExecutor
        if not force and app.current_executors() >= self._max_executors(app):
            return False
        conf = app.desc.get("conf", {})
        executor_id = app.get_next_executor_id()
        app.add_pending(executor_id)
        resources = dict(app.desc.get("resources") or {})
        opts: Dict[str, Any] = {
            "name": C.EXECUTOR_NAME_PREFIX + str(executor_id),
            "num_cpus": float(app.desc.get("ray_actor_cpu", app.desc["cores_per_executor"])),
            "memory": float(app.desc["memory_per_executor"]),
            "resources": resources,
            "max_restarts": -1, "max_task_retries": -1,
            "max_concurrency": int(app.desc["cores_per_executor"]) + 1,
        }
        pg_hex = conf.get(C.RAY_PLACEMENT_GROUP)
        if pg_hex:
            idxs = [int(x) for x in str(conf.get(C.RAY_BUNDLE_INDEXES, "")).split(",") if x != ""]
            opts["placement_group"] = pg_hex
            opts["placement_group_bundle_index"] = idxs[executor_id % len(idxs)] if idxs else -1
        handle = runtime.remote(SparkExecutor).options(**opts).remote(
            executor_id, app.id, self._name, int(app.desc["cores_per_executor"]), conf)
        app.by_actor[executor_id].handle = handle
        return True

    def register_executor(self, executor_id: int, node_ip: str, pid: int,
                          actor_executor_id: Optional[int] = None) -> bool:
        with self._lock:
            for app in self._apps.values():
                if executor_id in app.executors:
                    ok = app.register_executor(executor_id, node_ip, pid)
                    app.mark_started(executor_id)   # ExecutorStarted follows immediately
                    self._ensure_shuffle_service(app, node_ip)
                    return ok
        return False

    def _ensure_shuffle_service(self, app: ApplicationInfo, node_ip: str) -> None:
        """One external shuffle service actor per node, started with the first executor that
        registers from it (reference ``RayAppMaster.scala:151-167``)."""
        conf = app.desc.get("conf", {})
        if str(conf.get(C.SHUFFLE_SERVICE_ENABLED, "false")).lower() != "true":
            return
        if node_ip in self._shuffle_services:
            return
        from raydp_b200 import runtime
        from raydp_b200.spark.shuffle_service import RayExternalShuffleService, service_name
        handle = runtime.remote(RayExternalShuffleService).options(
            name=service_name(self._name, node_ip), num_cpus=0,
            resources={f"node:{node_ip}": 0.01}, max_concurrency=4).remote(node_ip, conf)
        self._shuffle_services[node_ip] = handle
        handle.start.remote()

    def shuffle_service_nodes(self) -> List[str]:
        with self._lock:
            return sorted(self._shuffle_services)

    def request_add_pending_restarted_executor(self, actor_executor_id: int) -> int:
        """A restarted executor actor asks for a fresh executor id but keeps its actor name."""
        with self._lock:
            for app in self._apps.values():
                ex = app.by_actor.get(actor_executor_id)
                if ex is None:
                    continue
                new_id = app.get_next_executor_id()
                app.executors.pop(ex.executor_id, None)
                app.restarted_executors[new_id] = ex.executor_id
                ex.executor_id, ex.registered, ex.started = new_id, False, False
                app.executors[new_id] = ex
                return new_id
        raise RayDPException(f"executor actor {actor_executor_id} is not known to the app master")

    def get_restarted_executors(self) -> Dict[int, int]:
        with self._lock:
            out: Dict[int, int] = {}
            for app in self._apps.values():
                out.update(app.restarted_executors)
            return out

    def registered_executor_actor_ids(self) -> List[int]:
        with self._lock:
            return sorted(a for app in self._apps.values() for a, ex in app.by_actor.items()
                          if ex.registered)

    def request_executors(self, app_id: str, requested_total: int) -> bool:
        """Dynamic allocation up-scaling (reference ``RayAppMaster.scala:173-180``)."""
        with self._lock:
            app = self._apps.get(app_id)
            if app is None:
                return False
            app.target_total = requested_total
            missing = requested_total - app.current_executors()
            for _ in range(max(0, missing)):
                if not self._request_new_executor(app):
                    break
            return True

    def kill_executors(self, app_id: str, actor_executor_ids: List[int]) -> bool:
        from raydp_b200 import runtime
        with self._lock:
            app = self._apps.get(app_id)
            if app is None:
                return False
            victims = [app.kill(i) for i in actor_executor_ids]
        for ex in victims:
            if ex is not None and ex.handle is not None:
                runtime.kill(ex.handle, no_restart=True)
        return True

    def unregister_application(self, app_id: str) -> bool:
        from raydp_b200 import runtime
        with self._lock:
            app = self._apps.pop(app_id, None)
        if app is None:
            return False
        app.state = "FINISHED"
        for ex in list(app.by_actor.values()):
            if ex.handle is not None:
                runtime.kill(ex.handle, no_restart=True)
        return True

    def application_info(self, app_id: str) -> Dict[str, Any]:
        with self._lock:
            app = self._apps.get(app_id)
            if app is None:
                return {}
            return {"id": app.id, "state": app.state, "target_total": app.target_total,
                    "executors": {e.executor_id: {"actor": e.actor_executor_id, "node_ip": e.node_ip,
                                                  "registered": e.registered, "pid": e.pid}
                                  for e in app.by_actor.values()},
                    "restarted": dict(app.restarted_executors)}

    # ------------------------------------------------------------------ driver agent
    def new_rdd_id(self) -> int:
        with self._lock:
            self._next_rdd_id += 1
            return self._next_rdd_id

    def pin_recoverable(self, rdd_id: int, plan) -> None:
        with self._lock:
            self._recoverable[rdd_id] = plan

    def get_recoverable_plan(self, rdd_id: int):
        with self._lock:
            return self._recoverable.get(rdd_id)

    def release_recoverable(self, rdd_id: int) -> bool:
        with self._lock:
            return self._recoverable.pop(rdd_id, None) is not None

    def persistent_rdd_ids(self) -> List[int]:
        with self._lock:
            return sorted(self._recoverable)

    # ------------------------------------------------------------------ stop
    def stop(self, cleanup_data: bool = True) -> None:
        from raydp_b200 import runtime
        for app_id in list(self._apps):
            self.unregister_application(app_id)
        self._recoverable.clear()
        for handle in self._shuffle_services.values():
            try:
                handle.stop.remote()
            except Exception:  # noqa: BLE001
                pass
        self._shuffle_services.clear()
        if cleanup_data:
            refs = [r for objs in self._objects.values() for r in objs]
            self._objects.clear()
            if refs:
                runtime.free(refs)
            runtime.exit_actor()
