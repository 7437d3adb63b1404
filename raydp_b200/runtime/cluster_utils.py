"""Local multi-node emulation with fault injection.

Parity: ``ray.cluster_utils.Cluster`` as used by reference ``test_spark_cluster.py:43-47,96,
280-292`` and ``test_spark_master_memory.py:9-19`` -- several virtual nodes on one host,
``add_node(num_cpus=, resources=)``, ``remove_node(node)`` (kills that node's processes and
loses the objects it stored), ``shutdown()``.
"""
from __future__ import annotations

from typing import Any, Dict, Optional

from raydp_b200 import runtime


class NodeHandle:
    def __init__(self, node_id: str, resources: Dict[str, float]):
        self.node_id = node_id
        self.resources = resources

    def __repr__(self):
        return f"Node({self.node_id}, {self.resources})"


class Cluster:
    def __init__(self, initialize_head: bool = False, connect: bool = False,
                 head_node_args: Optional[Dict[str, Any]] = None, **_ignored):
        self.head_node: Optional[NodeHandle] = None
        self.worker_nodes = []
        self._connected = False
        if initialize_head:
            args = dict(head_node_args or {})
            res = dict(args.get("resources") or {})
            runtime.init(address="local", num_cpus=args.get("num_cpus"),
                         num_gpus=args.get("num_gpus", 0), resources=res,
                         namespace=args.get("namespace"), _memory=args.get("memory"),
                         object_store_memory=args.get("object_store_memory"))
            self._connected = True
            self.head_node = NodeHandle(runtime.session_info()["node_id"], res)

    @property
    def address(self) -> str:
        return runtime.session_info().get("address", "")

    def connect(self, namespace: Optional[str] = None):
        if not runtime.is_initialized():
            runtime.init(namespace=namespace)
        self._connected = True
        return runtime.session_info()

    def add_node(self, num_cpus: float = 1, num_gpus: float = 0,
                 resources: Optional[Dict[str, float]] = None, memory: Optional[int] = None,
                 object_store_memory: Optional[int] = None, **_ignored) -> NodeHandle:
        res = {"CPU": float(num_cpus), "GPU": float(num_gpus)}
        if memory is not None:
            res["memory"] = float(memory)
        if object_store_memory is not None:
            res["object_store_memory"] = float(object_store_memory)
        res.update(resources or {})
        node_id = runtime._rpc("add_node", resources=res)
        node = NodeHandle(node_id, res)
        self.worker_nodes.append(node)
        return node

    def remove_node(self, node: NodeHandle, allow_graceful: bool = False) -> None:
        runtime._rpc("remove_node", node_id=node.node_id)
        if node in self.worker_nodes:
            self.worker_nodes.remove(node)

    def wait_for_nodes(self, timeout: float = 30) -> None:
        return None

    def shutdown(self) -> None:
        runtime.shutdown()
