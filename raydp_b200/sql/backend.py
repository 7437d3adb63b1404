"""Scheduler backends of the DataFrame engine: in-process threads or executor actors.

Parity: reference ``RayCoarseGrainedSchedulerBackend.scala:45-315`` -- the driver-side backend
that parses ``ray://<master>`` (``:81-117``), sends ``RegisterApplication`` with an
``ApplicationDescription`` and blocks until ``RegisteredApplication`` (``:123-194,239-263``),
forwards ``doRequestTotalExecutors`` / ``doKillExecutors`` (``:278-301``) and unregisters on
stop.  ``LocalBackend`` is the ``local[n]`` master used by unit tests (reference
``conftest.py:66-74``).
"""
from __future__ import annotations

import itertools
import threading
import time
from concurrent.futures import ThreadPoolExecutor
from typing import Any, Dict, List, Optional, Sequence

from raydp_b200.sql.executor import EXECUTOR_NAME_PREFIX, execute_action
from raydp_b200.sql.plan import Plan


class Backend:
    default_parallelism = 1

    def run(self, plan: Plan, action: str, args: dict, indices: Optional[Sequence[int]] = None) -> list:
        raise NotImplementedError

    def stop(self) -> None:
        pass


class LocalBackend(Backend):
    def __init__(self, threads: int = 2):
        self.default_parallelism = max(1, threads)
        self._pool = ThreadPoolExecutor(max_workers=self.default_parallelism,
                                        thread_name_prefix="rdp-local-exec")
        self._blocks: Dict = {}

    def run(self, plan, action, args, indices=None):
        idxs = list(range(plan.num_partitions)) if indices is None else list(indices)
        futs = [self._pool.submit(execute_action, plan, i, action, args, self._blocks) for i in idxs]
        return [f.result() for f in futs]

    def stop(self):
        self._pool.shutdown(wait=False)
        self._blocks.clear()


class ClusterBackend(Backend):
    """Talks to the ``<app>_SPARK_MASTER`` actor and fans partition tasks out to executors."""

    def __init__(self, master_url: str, app_name: str, conf: Dict[str, str]):
        from raydp_b200 import runtime
        self._rt = runtime
        self.master_name = parse_master_url(master_url)
        self.master = runtime.get_actor(self.master_name)
        self.conf = conf
        num = get_initial_target_executor_number(conf)
        cores = int(conf.get("spark.executor.cores", "1"))
        from raydp_b200.utils import parse_memory_size
        mem = parse_memory_size(conf.get("spark.executor.memory", "1g"))
        actor_cpu = conf.get("spark.ray.raydp_spark_executor.actor.resource.cpu",
                             conf.get("spark.ray.actor.resource.cpu", str(cores)))
        prefix = "spark.ray.raydp_spark_executor.actor.resource."
        custom = {k[len(prefix):]: float(v) for k, v in conf.items()
                  if k.startswith(prefix) and k[len(prefix):].lower() != "cpu"}
        desc = {"name": app_name, "num_executors": num, "cores_per_executor": cores,
                "memory_per_executor": mem, "ray_actor_cpu": float(actor_cpu),
                "resources": custom, "conf": dict(conf)}
        # RegisterApplication -> RegisteredApplication(app_id); executors come up asynchronously
        self.app_id = runtime.get(self.master.register_application.remote(desc))
        self._rr = itertools.count()
        self._lock = threading.Lock()
        self._executors: List[Any] = []
        self.default_parallelism = max(1, num * cores)
        self._wait_for_executors(min_count=max(1, num) if num > 0 else 0)

    def _wait_for_executors(self, min_count: int, timeout: float = 120.0) -> None:
        deadline = time.time() + timeout
        while True:
            ids = self._rt.get(self.master.registered_executor_actor_ids.remote())
            if len(ids) >= min_count:
                break
            if time.time() > deadline:
                raise TimeoutError(f"only {len(ids)}/{min_count} executors registered in {timeout}s")
            time.sleep(0.02)
        self._refresh(ids)

    def _refresh(self, ids=None) -> None:
        if ids is None:
            ids = self._rt.get(self.master.registered_executor_actor_ids.remote())
        with self._lock:
            self._executors = [self._rt.get_actor(EXECUTOR_NAME_PREFIX + str(i)) for i in sorted(ids)]

    def executors(self) -> List[Any]:
        if not self._executors:
            self._refresh()
        return list(self._executors)

    def executor_actor_ids(self) -> List[int]:
        return sorted(self._rt.get(self.master.registered_executor_actor_ids.remote()))

    def run(self, plan, action, args, indices=None, preferred: Optional[Dict[int, int]] = None):
        idxs = list(range(plan.num_partitions)) if indices is None else list(indices)
        execs = self.executors()
        if not execs:
            raise RuntimeError("no executors are registered")
        plan_ref = self._rt.put(plan)
        try:
            refs = []
            for i in idxs:
                slot = (preferred or {}).get(i, next(self._rr)) % len(execs)
                refs.append(execs[slot].run_partition.remote(plan_ref, i, action, args))
            return self._rt.get(refs)
        finally:
            self._rt.free([plan_ref])

    # dynamic allocation ---------------------------------------------------
    def request_total_executors(self, total: int) -> bool:
        ok = self._rt.get(self.master.request_executors.remote(self.app_id, total))
        self._wait_for_executors(min_count=total)
        return ok

    def kill_executors(self, actor_ids: Sequence[int]) -> bool:
        ok = self._rt.get(self.master.kill_executors.remote(self.app_id, list(actor_ids)))
        self._refresh()
        return ok

    def stop(self):
        try:
            self._rt.get(self.master.unregister_application.remote(self.app_id), timeout=30)
        except Exception:  # noqa: BLE001 - master already gone
            pass


def parse_master_url(url: str) -> str:
    """``ray://<master actor name>`` -> actor name (reference ``:81-117`` parses
    ``ray://name@host:port``; there is no host/port because actors are looked up by name)."""
    if not url.startswith("ray"):
        raise ValueError(f"Invalid master URL: {url}")
    rest = url[len("ray://"):] if url.startswith("ray://") else ""
    return rest.split("@")[0] if rest else ""


def get_initial_target_executor_number(conf: Dict[str, str], num_executors: int = 2) -> int:
    """Initial executor count under static or dynamic allocation.

    Parity: Spark's ``SchedulerBackendUtils.getInitialTargetExecutorNumber`` as unit-tested by
    the reference's only JVM test (``TestRayCoarseGrainedSchedulerBackend.java:32-65``).
    """
    dyn = str(conf.get("spark.dynamicAllocation.enabled", "false")).lower() == "true"
    if dyn:
        lo = int(conf.get("spark.dynamicAllocation.minExecutors", "0"))
        init = int(conf.get("spark.dynamicAllocation.initialExecutors", str(lo)))
        hi = int(conf.get("spark.dynamicAllocation.maxExecutors", str(2 ** 31 - 1)))
        inst = int(conf.get("spark.executor.instances", "0"))
        target = max(lo, init, inst)
        if not lo <= target <= hi:
            raise ValueError(f"initial executor number {target} must between min executor number "
                             f"{lo} and max executor number {hi}")
        return target
    n = int(conf.get("spark.executor.instances", str(num_executors)))
    return n if n >= 0 else num_executors
