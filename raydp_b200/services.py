"""Service interfaces of a cluster that runs on the actor runtime.

The names are the reference's extension points (``python/raydp/services.py:22-107``): ``Cluster`` --
subclassed by :class:`raydp_b200.spark.ray_cluster.SparkCluster` -- plus ``ClusterMaster`` and
``ClusterWorker`` for the two kinds of service processes.  Here the two service kinds share one
base (:class:`Service`), and ``Cluster.add_worker`` keeps count of the workers it started and
stops the whole cluster when starting one fails, so a half-built cluster never lingers.
"""
from __future__ import annotations

import abc
from typing import Any, Dict, Optional

Resources = Dict[str, float]


class Service(abc.ABC):
    """A long-running process of the cluster, addressable by host."""

    @abc.abstractmethod
    def get_host(self) -> str:
        ...

    @abc.abstractmethod
    def stop(self) -> None:
        ...


class ClusterMaster(Service):
    @abc.abstractmethod
    def start_up(self, resources: Optional[Resources] = None) -> str:
        """Bring the master up; the return value is its URL."""

    @abc.abstractmethod
    def get_master_url(self) -> str:
        ...


class ClusterWorker(Service):
    @abc.abstractmethod
    def start_up(self) -> Optional[str]:
        ...


class Cluster(abc.ABC):
    """One master, workers added on demand."""

    def __init__(self, master_resources_requirement: Optional[Resources] = None):
        self._master_resources_requirement = master_resources_requirement
        self._num_workers = 0

    @property
    def num_workers(self) -> int:
        return self._num_workers

    def add_worker(self, resources_requirement: Resources, **kwargs: Any):
        try:
            worker = self._set_up_worker(resources_requirement, kwargs)
        except BaseException:
            self.stop()
            raise
        self._num_workers += 1
        return worker

    @abc.abstractmethod
    def _set_up_master(self, resources: Optional[Resources], kwargs: Optional[Dict[Any, Any]]):
        """Start the master and block until it answers."""

    @abc.abstractmethod
    def _set_up_worker(self, resources: Resources, kwargs: Dict[str, Any]):
        """Start one worker holding ``resources``."""

    @abc.abstractmethod
    def get_cluster_url(self) -> str:
        """What clients connect to, e.g. ``ray://<master actor name>``."""

    @abc.abstractmethod
    def stop(self, cleanup_data: bool = True) -> None:
        """Stop workers, then the master."""
