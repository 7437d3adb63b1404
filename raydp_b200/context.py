"""Session lifecycle: ``init_spark`` / ``stop_spark``.

Behaviour kept from the reference (``python/raydp/context.py``): one session per process behind a
re-entrant lock (``:178-179``); the runtime is started on demand (``:215-217``); a second
``init_spark`` raises ``"The spark environment has inited."`` (``:240``); a failed start is torn down
(``:234-238``); a placement-group *strategy* creates one ``{"CPU": cores, "memory": bytes}`` bundle per
executor, waits for it and removes it again at stop, while an existing group is only referenced
(``:119-140``) -- either way the executors learn about it through ``spark.ray.placement_group`` /
``spark.ray.bundle_indexes``; ``stop_spark(cleanup_data=False)`` leaves the master actor, and the
datasets it owns, alive (``:243-254``); ``stop_spark`` runs at interpreter exit (``:257``).

The shape is different: the request is a frozen :class:`SessionRequest`, the live pieces (placement
group, cluster, session) sit in one :class:`_LiveSession` that is built by ``start()`` and unwound by
``shutdown()`` in reverse order, so a failure half way through start-up releases exactly what exists.
"""
from __future__ import annotations

import atexit
import threading
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Union

from raydp_b200 import runtime
from raydp_b200.spark import configs as C
from raydp_b200.spark.ray_cluster import SparkCluster
from raydp_b200.sql.session import SparkSession, reset_builder_singletons
from raydp_b200.utils import parse_memory_size

ALREADY_STARTED = "The spark environment has inited."


@dataclass(frozen=True)
class SessionRequest:
    app_name: str
    num_executors: int
    executor_cores: int
    executor_memory: Union[str, int]
    enable_hive: bool = False
    fault_tolerant_mode: bool = True
    placement_group_strategy: Optional[str] = None
    placement_group: object = None
    placement_group_bundle_indexes: Optional[List[int]] = None
    configs: Dict[str, str] = field(default_factory=dict)

    def memory_bytes(self) -> int:
        mem = self.executor_memory
        return parse_memory_size(mem) if isinstance(mem, str) else int(mem)


class _LiveSession:
    """Everything ``init_spark`` brought up, in creation order."""

    def __init__(self, request: SessionRequest):
        self.request = request
        self.configs: Dict[str, str] = dict(request.configs)
        self.owned_group = None            # a placement group made here (strategy given) -> removed at stop
        self.cluster: Optional[SparkCluster] = None
        self.session: Optional[SparkSession] = None

    # ---- start-up
    def _reserve(self) -> None:
        req = self.request
        group, indexes = req.placement_group, req.placement_group_bundle_indexes
        if req.placement_group_strategy is not None:
            bundle = {"CPU": req.executor_cores, "memory": req.memory_bytes()}
            group = runtime.placement_group([dict(bundle) for _ in range(req.num_executors)],
                                            strategy=req.placement_group_strategy)
            runtime.get(group.ready())
            self.owned_group, indexes = group, None
        if group is None:
            return
        if indexes is None:
            indexes = range(group.bundle_count)
        self.configs[C.RAY_PLACEMENT_GROUP] = group.id.hex()
        self.configs[C.RAY_BUNDLE_INDEXES] = ",".join(str(i) for i in indexes)

    def start(self) -> SparkSession:
        req = self.request
        self._reserve()
        self.cluster = SparkCluster(req.app_name, req.num_executors, req.executor_cores, req.executor_memory,
                                    req.enable_hive, self.configs)
        self.session = self.cluster.get_spark_session()
        if req.fault_tolerant_mode:
            self.cluster.connect_spark_driver_to_ray()
        return self.session

    # ---- tear-down (reverse order; every step tolerates "was never created")
    def shutdown(self, cleanup_data: bool) -> bool:
        """Returns True when nothing is left alive (the master actor survives ``cleanup_data=False``)."""
        if self.session is not None:
            self.session.stop()
            self.session = None
            reset_builder_singletons()
        if self.cluster is not None:
            self.cluster.stop(cleanup_data)
            if cleanup_data:
                self.cluster = None
        if self.owned_group is not None:
            runtime.remove_placement_group(self.owned_group)
            self.owned_group = None
        return self.cluster is None


_lock = threading.RLock()
_live: Optional[_LiveSession] = None


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
    global _live
    request = SessionRequest(app_name, num_executors, executor_cores, executor_memory, enable_hive,
                             fault_tolerant_mode, placement_group_strategy, placement_group,
                             placement_group_bundle_indexes, dict(configs or {}))
    if not runtime.is_initialized():
        runtime.init()
    with _lock:
        if _live is not None:
            raise Exception(ALREADY_STARTED)
        candidate = _LiveSession(request)
        try:
            session = candidate.start()
        except BaseException:
            try:
                candidate.shutdown(cleanup_data=True)
            except Exception:  # noqa: BLE001  (the original error is the one worth reporting)
                pass
            raise
        _live = candidate
        return session


def stop_spark(cleanup_data: bool = True) -> None:
    """Stop the Spark session and executors.  With ``cleanup_data=False`` the master actor and
    the converted datasets it holds stay alive until the next ``stop_spark(True)``."""
    global _live
    with _lock:
        if _live is None:
            return
        gone = True
        try:
            if runtime.is_initialized():
                gone = _live.shutdown(cleanup_data)
        finally:
            if gone or not runtime.is_initialized():
                _live = None


atexit.register(stop_spark)
