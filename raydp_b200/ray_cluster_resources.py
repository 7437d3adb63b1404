"""Cluster resource probe.

Parity: reference ``python/raydp/ray_cluster_resources.py:25-79`` (``ClusterResources.total_alive_nodes``
/ ``satisfy(request)``); the reference version imports a removed Ray module and is dead code there --
this one works against the built-in runtime.
"""
from typing import Dict, List

from raydp_b200 import runtime


class ClusterResources:
    @classmethod
    def total_alive_nodes(cls) -> int:
        return sum(1 for n in runtime.nodes() if n["Alive"])

    @classmethod
    def satisfy(cls, request: Dict[str, float]) -> List[str]:
        """Node ids whose *total* resources cover ``request``."""
        out = []
        for node in runtime.nodes():
            if node["Alive"] and all(node["Resources"].get(k, 0.0) >= v for k, v in request.items()):
                out.append(node["NodeID"])
        return out
