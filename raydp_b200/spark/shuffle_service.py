"""Per-node external shuffle service actor.

Parity: reference ``RayExternalShuffleService.scala:26-57`` + ``ExternalShuffleServiceUtils.java:25-42``
-- when ``spark.shuffle.service.enabled`` is true the app master lazily starts one service actor
per node (pinned with a ``node:<ip>`` resource of 0.01, ``RayAppMaster.scala:151-167``) so that
shuffle output survives executors being killed by dynamic allocation.

Here shuffle output is a set of sealed shared-memory objects, so "serving" it needs no data path:
the service *owns* the blocks written by the executors on its node (they are put with the
service as owner), keeps them registered per shuffle id, and frees them when the shuffle or the
application ends -- independently of the lifetime of the executor that wrote them.
"""
from __future__ import annotations

import threading
from typing import Any, Dict, List

SERVICE_NAME_PREFIX = "raydp-shuffle-service-"


def service_name(app_name: str, node_ip: str) -> str:
    return f"{SERVICE_NAME_PREFIX}{app_name}-{node_ip}"


class RayExternalShuffleService:
    def __init__(self, node_ip: str, conf: Dict[str, str]):
        self.node_ip = node_ip
        # only shuffle-related settings matter to the service (reference getShuffleConf)
        self.conf = {k: v for k, v in (conf or {}).items()
                     if k.startswith("spark.shuffle") or k == "spark.local.dir"}
        self._blocks: Dict[str, List[Any]] = {}
        self._lock = threading.Lock()

    def start(self) -> str:
        return self.node_ip

    def register_blocks(self, shuffle_id: str, refs: List[Any]) -> int:
        with self._lock:
            held = self._blocks.setdefault(shuffle_id, [])
            held.extend(refs)
            return len(held)

    def block_count(self, shuffle_id: str = "") -> int:
        with self._lock:
            if shuffle_id:
                return len(self._blocks.get(shuffle_id, []))
            return sum(len(v) for v in self._blocks.values())

    def remove_shuffle(self, shuffle_id: str) -> int:
        from raydp_b200 import runtime
        with self._lock:
            refs = self._blocks.pop(shuffle_id, [])
        if refs:
            runtime.free(refs)
        return len(refs)

    def stop(self) -> None:
        from raydp_b200 import runtime
        with self._lock:
            refs = [r for v in self._blocks.values() for r in v]
            self._blocks.clear()
        if refs:
            runtime.free(refs)
        runtime.exit_actor()
