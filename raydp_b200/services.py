"""Cluster service ABCs.

Parity: reference ``python/raydp/services.py:22-107`` -- ``Cluster`` is the only one that is
subclassed (by ``SparkCluster``); ``ClusterMaster`` / ``ClusterWorker`` are kept as the same
extension points.
"""
from abc import ABC, abstractmethod
from typing import Any, Dict, Optional


class Cluster(ABC):
    """A cluster = one master plus workers added on demand."""

    def __init__(self, master_resources_requirement: Optional[Dict[str, float]] = None):
        self._master_resources_requirement = master_resources_requirement

    @abstractmethod
    def _set_up_master(self, resources: Optional[Dict[str, float]], kwargs: Dict[Any, Any]):
        """Start the master and block until it is ready."""

    def add_worker(self, resources_requirement: Dict[str, float], **kwargs):
        return self._set_up_worker(resources_requirement, kwargs)

    @abstractmethod
    def _set_up_worker(self, resources: Dict[str, float], kwargs: Dict[str, str]):
        """Start one worker with the given resources."""

    @abstractmethod
    def get_cluster_url(self) -> str:
        """URL that clients use to connect to this cluster."""

    @abstractmethod
    def stop(self, cleanup_data: bool = True):
        """Stop workers and master."""


class ClusterMaster(ABC):
    @abstractmethod
    def start_up(self, resources: Optional[Dict[str, float]] = None) -> str:
        """Start the master; returns its URL."""

    @abstractmethod
    def get_master_url(self) -> str:
        pass

    @abstractmethod
    def get_host(self) -> str:
        pass

    @abstractmethod
    def stop(self):
        pass


class ClusterWorker(ABC):
    @abstractmethod
    def start_up(self) -> Optional[str]:
        pass

    @abstractmethod
    def get_host(self) -> str:
        pass

    @abstractmethod
    def stop(self):
        pass
