"""Built-in actor runtime: the Ray surface the reference uses, with no Ray dependency.

Parity map: ``init/shutdown/is_initialized`` (reference
``context.py:215-217``), ``@remote`` classes/functions with ``.options(name=, num_cpus=,
memory=, resources=, placement_group=, placement_group_bundle_index=, max_restarts=,
max_task_retries=, max_concurrency=, max_retries=)`` (reference ``spark/ray_cluster.py:69-78``,
``spark/dataset.py:54,211``, ``RayExecutorUtils.java:46-61``), ``get/put/wait``,
``get_actor`` and ``actor.exit_actor()`` (reference ``ray_cluster_master.py:247``),
``available_resources`` (reference ``test_from_spark.py:54-66``), placement groups
(reference ``context.py:128-140``), node IPs and the runtime context.
"""
from __future__ import annotations

import atexit
import functools
import inspect
import os
import threading
from typing import Any, Dict, List, Optional, Sequence, Tuple

from raydp_b200.runtime import protocol as P
from raydp_b200.runtime.protocol import (ActorDiedError, GetTimeoutError, ObjectLostError,
                                         RuntimeErrorBase, TaskError)

__all__ = [
    "init", "shutdown", "is_initialized", "remote", "get", "put", "wait", "free", "get_actor",
    "kill", "available_resources", "cluster_resources", "nodes", "get_node_ip_address",
    "get_runtime_context", "ObjectRef", "ActorHandle", "placement_group",
    "remove_placement_group", "placement_group_table", "get_placement_group", "list_actors",
    "TaskError", "ActorDiedError", "GetTimeoutError", "ObjectLostError", "exit_actor", "worker_logs",
]

_state_lock = threading.RLock()
_head = None           # Head instance when this process is the driver
_client = None         # RpcClient or _LocalClient
_info: Dict[str, Any] = {}
_actor_ctx: Dict[str, Any] = {}


class _ActorExit(BaseException):
    pass


class _LocalClient:
    """Driver-side shortcut: call the in-process head without a socket hop."""

    def __init__(self, head):
        self._head = head
        self._info = {"node_id": head.head_node, "worker_id": "driver", "namespace": head.namespace}

    def request(self, op: str, **kwargs):
        return self._head.handle(op, kwargs, self._info)

    def close(self):
        pass


# --------------------------------------------------------------------------- lifecycle
def init(address: Optional[str] = None, num_cpus: Optional[int] = None,
         num_gpus: Optional[int] = None, resources: Optional[Dict[str, float]] = None,
         namespace: Optional[str] = None, ignore_reinit_error: bool = False,
         include_dashboard: bool = False, object_store_memory: Optional[int] = None,
         _memory: Optional[int] = None, log_to_driver: Optional[bool] = None, **_ignored):
    """Start (``address in (None, "local")``) or join (``"auto"`` / worker env) a runtime."""
    global _head, _client, _info
    with _state_lock:
        if _client is not None:
            # "connecting" to the runtime this process already hosts (``Cluster(...)`` followed by
            # ``init(address=cluster.address)``) is a no-op
            if ignore_reinit_error or address == "auto" or (address and address == _info.get("address")):
                return _info
            raise RuntimeError("Maybe you called runtime.init twice by accident? "
                               "pass ignore_reinit_error=True to allow it")
        if os.environ.get(P.ENV_HEAD) and address in (None, "auto"):
            _connect_from_env()
            return _info
        from raydp_b200.runtime.head import Head
        _head = Head(num_cpus=num_cpus, num_gpus=num_gpus, resources=resources,
                     namespace=namespace, object_store_memory=object_store_memory,
                     memory=_memory, log_to_driver=log_to_driver)
        _client = _LocalClient(_head)
        _info = _client.request("session_info")
        _info.update({"node_id": _head.head_node, "node_ip": _info["head_node_ip"],
                      "is_driver": True})
        return _info


def _connect_from_env() -> None:
    global _client, _info
    with _state_lock:
        if _client is not None:
            return
        head = os.environ[P.ENV_HEAD]
        auth = bytes.fromhex(os.environ[P.ENV_AUTH])
        client = P.RpcClient(head, auth)
        # hello goes through the raw connection before any request
        with client._send_lock:
            client._conn.send(("client", os.environ[P.ENV_NODE], os.environ.get(P.ENV_WORKER, "?"),
                               os.environ.get(P.ENV_NAMESPACE)))
        _client = client
        _info = client.request("session_info")
        _info.update({"node_id": os.environ[P.ENV_NODE], "node_ip": os.environ[P.ENV_NODE_IP],
                      "is_driver": False})
        if os.environ.get(P.ENV_NAMESPACE):
            _info["namespace"] = os.environ[P.ENV_NAMESPACE]


def is_initialized() -> bool:
    return _client is not None


def shutdown() -> None:
    global _head, _client, _info
    with _state_lock:
        if _client is None:
            return
        try:
            if _head is not None:
                _head.shutdown()
            else:
                _client.close()
        finally:
            _head = None
            _client = None
            _info = {}


def _register_exit_hooks() -> None:
    """Shut the head down as early as possible at interpreter exit.  ``threading._shutdown`` runs
    BEFORE ``atexit`` and first disables every ``ThreadPoolExecutor`` (concurrent.futures'
    ``_python_exit``): from then on the head cannot schedule the handshake of a worker process that is
    still starting, and its launch thread would sit in ``ready.wait(60)`` while the interpreter waits
    for that thread.  Callbacks registered with ``threading._register_atexit`` run in reverse order of
    registration, i.e. ours (registered after concurrent.futures was imported) runs first."""
    import threading
    import concurrent.futures.thread  # noqa: F401  (its own exit hook must be registered BEFORE ours)
    try:
        threading._register_atexit(shutdown)          # CPython >= 3.9
    except Exception:  # noqa: BLE001
        pass
    atexit.register(shutdown)


_register_exit_hooks()


def _rpc(op: str, **kwargs):
    if _client is None:
        raise RuntimeErrorBase("runtime is not initialised; call raydp_b200.runtime.init()")
    return _client.request(op, **kwargs)


def _auto_init() -> None:
    if _client is None:
        with _state_lock:
            if _client is None:
                init(ignore_reinit_error=True)


# --------------------------------------------------------------------------- objects
class ObjectRef:
    """Handle to a value in the object directory (future + location-transparent pointer)."""

    __slots__ = ("_id", "_owner", "__weakref__")

    def __init__(self, oid: str, owner: Optional[str] = None):
        self._id = oid
        self._owner = owner

    def hex(self) -> str:
        return self._id

    def binary(self) -> bytes:
        return bytes.fromhex(self._id)

    def __reduce__(self):
        return (ObjectRef, (self._id, self._owner))

    def __hash__(self):
        return hash(self._id)

    def __eq__(self, other):
        return isinstance(other, ObjectRef) and other._id == self._id

    def __repr__(self):
        return f"ObjectRef({self._id})"


def put(value: Any, _owner: Optional["ActorHandle"] = None) -> ObjectRef:
    _auto_init()
    oid = P.new_id()
    desc = P.make_descriptor(value, _info["session"], _info["node_id"], oid)
    owner_actor = _owner._actor_id if _owner is not None else _actor_ctx.get("actor_id")
    _rpc("put", oid=oid, desc=desc, node_id=_info["node_id"], owner_actor=owner_actor)
    return ObjectRef(oid, owner_actor)


def _get_one(ref: ObjectRef, timeout: Optional[float], pin: bool = False):
    for _attempt in range(8):
        desc = _rpc("get", oid=ref._id, timeout=timeout)
        try:
            return P.read_descriptor(desc, pin=pin)
        except ObjectLostError:
            # the segment vanished under us (node removed): ask for lineage reconstruction
            _rpc("report_lost", oid=ref._id)
    raise ObjectLostError(f"object {ref._id} could not be recovered")


def get(refs, timeout: Optional[float] = None, _pin: bool = False):
    if isinstance(refs, ObjectRef):
        return _get_one(refs, timeout, _pin)
    if isinstance(refs, (list, tuple)):
        return [_get_one(r, timeout, _pin) for r in refs]
    raise TypeError(f"get() expects an ObjectRef or a list of them, got {type(refs)}")


def wait(refs: Sequence[ObjectRef], num_returns: int = 1, timeout: Optional[float] = None
         ) -> Tuple[List[ObjectRef], List[ObjectRef]]:
    by_id = {r._id: r for r in refs}
    ready, rest = _rpc("wait", oids=[r._id for r in refs], num_returns=num_returns, timeout=timeout)
    return [by_id[i] for i in ready], [by_id[i] for i in rest]


def free(refs: Sequence[ObjectRef]) -> None:
    if _client is None:
        return
    _rpc("free", oids=[r._id for r in refs])


def object_locations(refs: Sequence[ObjectRef]) -> List[Optional[str]]:
    return _rpc("object_locations", oids=[r._id for r in refs])


# --------------------------------------------------------------------------- options
def _normalise_resources(opts: Dict[str, Any], default_cpus: float) -> Dict[str, float]:
    res: Dict[str, float] = {}
    for k, v in (opts.get("resources") or {}).items():
        res[k] = float(v)
    cpus = opts.get("num_cpus")
    if cpus is not None:
        res["CPU"] = float(cpus)
    elif "CPU" not in res:
        res["CPU"] = float(default_cpus)
    if opts.get("num_gpus") is not None:
        res["GPU"] = float(opts["num_gpus"])
    if opts.get("memory") is not None:
        res["memory"] = float(opts["memory"])
    return {k: v for k, v in res.items() if v > 0}


def _pg_fields(opts: Dict[str, Any]) -> Tuple[Optional[str], int]:
    pg = opts.get("placement_group")
    strategy = opts.get("scheduling_strategy")
    bundle = opts.get("placement_group_bundle_index", -1)
    if strategy is not None and hasattr(strategy, "placement_group"):
        pg = strategy.placement_group
        bundle = getattr(strategy, "placement_group_bundle_index", -1)
    if pg is None or pg == "default":
        return None, -1
    return (pg.id.hex() if isinstance(pg, PlacementGroup) else str(pg)), int(bundle if bundle is not None else -1)


class PlacementGroupSchedulingStrategy:
    def __init__(self, placement_group, placement_group_bundle_index: int = -1,
                 placement_group_capture_child_tasks: Optional[bool] = None):
        self.placement_group = placement_group
        self.placement_group_bundle_index = placement_group_bundle_index


# --------------------------------------------------------------------------- tasks
class RemoteFunction:
    def __init__(self, func, opts: Dict[str, Any]):
        self._func = func
        self._opts = dict(opts)
        self._blob: Optional[bytes] = None
        functools.update_wrapper(self, func)

    def options(self, **opts) -> "RemoteFunction":
        merged = dict(self._opts)
        merged.update(opts)
        clone = RemoteFunction(self._func, merged)
        clone._blob = self._blob
        return clone

    def remote(self, *args, **kwargs) -> ObjectRef:
        _auto_init()
        if self._blob is None:
            self._blob = P.dumps(self._func)
        pg, bundle = _pg_fields(self._opts)
        oid = _rpc("submit_task", func_blob=self._blob, args_blob=P.dumps((args, kwargs)),
                   resources=_normalise_resources(self._opts, 1.0),
                   max_retries=int(self._opts.get("max_retries", 3)),
                   name=getattr(self._func, "__name__", "task"),
                   placement_group=pg, bundle_index=bundle,
                   prefer_ip=self._opts.get("_prefer_node_ip"))
        return ObjectRef(oid)

    def __call__(self, *args, **kwargs):
        raise TypeError("Remote functions cannot be called directly; use .remote()")


# --------------------------------------------------------------------------- actors
class _ActorMethod:
    def __init__(self, handle: "ActorHandle", name: str):
        self._handle = handle
        self._name = name

    def remote(self, *args, **kwargs) -> ObjectRef:
        oid = _rpc("call_actor", actor_id=self._handle._actor_id, method=self._name,
                   args_blob=P.dumps((args, kwargs)))
        return ObjectRef(oid, self._handle._actor_id)

    def options(self, **_):
        return self


class ActorHandle:
    def __init__(self, actor_id: str, class_name: str = "", create_oid: Optional[str] = None):
        self._actor_id = _Id(actor_id)
        self._class_name = class_name
        self._create_oid = create_oid

    def __getattr__(self, name: str):
        if name.startswith("_"):
            raise AttributeError(name)
        return _ActorMethod(self, name)

    def __reduce__(self):
        return (ActorHandle, (self._actor_id, self._class_name, self._create_oid))

    def __repr__(self):
        return f"Actor({self._class_name}, {self._actor_id[:12]})"

    @property
    def _ready_ref(self) -> ObjectRef:
        return ObjectRef(self._create_oid)

    def _wait_ready(self, timeout: Optional[float] = None) -> "ActorHandle":
        get(ObjectRef(self._create_oid), timeout=timeout)
        return self

    def _state(self) -> str:
        return _rpc("actor_state", actor_id=self._actor_id)


class ActorClass:
    def __init__(self, cls, opts: Dict[str, Any]):
        self._cls = cls
        self._opts = dict(opts)
        self._blob: Optional[bytes] = None
        self.__name__ = getattr(cls, "__name__", "Actor")
        self.__doc__ = cls.__doc__

    def options(self, **opts) -> "ActorClass":
        merged = dict(self._opts)
        merged.update(opts)
        clone = ActorClass(self._cls, merged)
        clone._blob = self._blob
        return clone

    def remote(self, *args, **kwargs) -> ActorHandle:
        _auto_init()
        if self._blob is None:
            self._blob = P.dumps(self._cls)
        o = self._opts
        pg, bundle = _pg_fields(o)
        lifetime_ns = o.get("namespace") or _info.get("namespace")
        opts = {"name": o.get("name"), "namespace": lifetime_ns,
                "resources": _normalise_resources(o, 0.0),
                "max_restarts": int(o.get("max_restarts", 0)),
                "max_task_retries": int(o.get("max_task_retries", 0)),
                "max_concurrency": int(o.get("max_concurrency", 1)),
                "placement_group": pg, "bundle_index": bundle,
                "get_if_exists": bool(o.get("get_if_exists", False)),
                "env": (o.get("runtime_env") or {}).get("env_vars")}
        actor_id, create_oid = _rpc("create_actor", cls_blob=self._blob,
                                    args_blob=P.dumps((args, kwargs)),
                                    class_name=self.__name__, opts=opts)
        return ActorHandle(actor_id, self.__name__, create_oid)

    def __call__(self, *args, **kwargs):
        raise TypeError("Actors cannot be instantiated directly; use .remote()")


def remote(*args, **opts):
    """``@remote`` / ``@remote(num_cpus=..)`` for functions and classes."""
    def wrap(target):
        if inspect.isclass(target):
            return ActorClass(target, opts)
        return RemoteFunction(target, opts)
    if len(args) == 1 and not opts and (inspect.isclass(args[0]) or callable(args[0])):
        return wrap(args[0])
    return wrap


def get_actor(name: str, namespace: Optional[str] = None) -> ActorHandle:
    _auto_init()
    actor_id, class_name, create_oid = _rpc("get_actor", name=name,
                                            namespace=namespace or _info.get("namespace"))
    return ActorHandle(actor_id, class_name, create_oid)


def kill(actor: ActorHandle, no_restart: bool = True) -> None:
    _rpc("kill_actor", actor_id=actor._actor_id, no_restart=no_restart)


def exit_actor() -> None:
    """Terminate the current actor process intentionally (no restart)."""
    if not _actor_ctx:
        raise RuntimeError("exit_actor() called outside an actor")
    raise _ActorExit()


class _Id(str):
    """Identifier string that also answers ``.hex()`` like the binary ids of the reference's runtime."""

    def hex(self) -> str:
        return str(self)


class ActorState(dict):
    """One ``list_actors`` record: a dict whose keys also read as attributes (``state.required_resources``)."""

    def __getattr__(self, name: str):
        try:
            return self[name]
        except KeyError as exc:
            raise AttributeError(name) from exc


def list_actors(filters: Optional[List[Tuple[str, str, Any]]] = None, detail: bool = False,
                **_ignored) -> List[ActorState]:
    filters = [(k, op, str(v) if k == "actor_id" else v) for k, op, v in (filters or [])]
    return [ActorState(r) for r in _rpc("list_actors", filters=filters)]


# --------------------------------------------------------------------------- cluster info
def available_resources() -> Dict[str, float]:
    return _rpc("resources", which="available")


def cluster_resources() -> Dict[str, float]:
    return _rpc("resources", which="total")


def nodes() -> List[dict]:
    return _rpc("nodes")


def worker_logs(name: Optional[str] = None, tail_bytes: int = 16384) -> Dict[str, dict]:
    """Tails of the per-worker log files of this session (stdout / stderr of every actor and task worker; each
    file starts with a ``:job_id:<id>`` header, as the reference's Java agent writes it).  ``name`` filters on the
    worker tag (``actor-<name>`` / ``task``).  ``init(log_to_driver=True)`` or ``RAYDP_B200_LOG_TO_DRIVER=1``
    additionally echoes the files on the driver's stderr while the workers run."""
    return _rpc("worker_logs", name=name, tail_bytes=tail_bytes)


def get_node_ip_address() -> str:
    return _info.get("node_ip", "127.0.0.1")


def session_info() -> Dict[str, Any]:
    return dict(_info)


class RuntimeContext:
    @property
    def namespace(self) -> str:
        return _info.get("namespace", "")

    @property
    def job_id(self) -> str:
        return _info.get("job_id", "")

    def get_job_id(self) -> str:
        return self.job_id

    def get_node_id(self) -> str:
        return _info.get("node_id", "")

    def get_actor_id(self) -> Optional[str]:
        return _actor_ctx.get("actor_id")

    @property
    def actor_id(self) -> Optional[str]:
        return _actor_ctx.get("actor_id")

    def get_actor_name(self) -> Optional[str]:
        return _actor_ctx.get("name")

    @property
    def was_current_actor_reconstructed(self) -> bool:
        return bool(_actor_ctx.get("was_restarted", False))

    def get_assigned_gpus(self) -> List[int]:
        return list(_actor_ctx.get("gpus", []))

    def get(self) -> Dict[str, Any]:
        return {"namespace": self.namespace, "job_id": self.job_id,
                "node_id": self.get_node_id(), "actor_id": self.get_actor_id()}


def get_runtime_context() -> RuntimeContext:
    return RuntimeContext()


def _set_actor_context(ctx: Dict[str, Any]) -> None:
    _actor_ctx.clear()
    _actor_ctx.update(ctx)


# --------------------------------------------------------------------------- placement groups
class _PGId:
    def __init__(self, hexid: str):
        self._hex = hexid

    def hex(self) -> str:
        return self._hex

    def __eq__(self, other):
        return isinstance(other, _PGId) and other._hex == self._hex

    def __hash__(self):
        return hash(self._hex)

    def __repr__(self):
        return f"PlacementGroupID({self._hex})"


class PlacementGroup:
    def __init__(self, pg_id: str, bundles: List[Dict[str, float]], ready_oid: Optional[str] = None):
        self.id = _PGId(pg_id)
        self.bundle_specs = [dict(b) for b in bundles]
        self._ready_oid = ready_oid

    @property
    def bundle_count(self) -> int:
        return len(self.bundle_specs)

    def ready(self) -> ObjectRef:
        return ObjectRef(self._ready_oid)

    def wait(self, timeout_seconds: float = 30) -> bool:
        try:
            get(self.ready(), timeout=timeout_seconds)
            return True
        except GetTimeoutError:
            return False

    def __reduce__(self):
        return (PlacementGroup, (self.id.hex(), self.bundle_specs, self._ready_oid))

    def __eq__(self, other):
        return isinstance(other, PlacementGroup) and other.id == self.id

    def __hash__(self):
        return hash(self.id)


def placement_group(bundles: List[Dict[str, float]], strategy: str = "PACK", name: str = "",
                    lifetime: Optional[str] = None) -> PlacementGroup:
    _auto_init()
    pg_id, ready_oid = _rpc("create_pg", bundles=[{k: float(v) for k, v in b.items()} for b in bundles],
                            strategy=strategy, name=name)
    return PlacementGroup(pg_id, bundles, ready_oid)


def remove_placement_group(pg: PlacementGroup) -> None:
    _rpc("remove_pg", pg_id=pg.id.hex())


def placement_group_table(pg: Optional[PlacementGroup] = None) -> Dict[str, Any]:
    return _rpc("pg_table", pg_id=pg.id.hex() if pg is not None else None)


def get_placement_group(name_or_hex: str) -> PlacementGroup:
    table = _rpc("pg_table", pg_id=None)
    for pg_id, row in table.items():
        if pg_id == name_or_hex or (row["name"] and row["name"] == name_or_hex):
            if row["state"] == "REMOVED":
                break
            bundles = [row["bundles"][i] for i in sorted(row["bundles"])]
            return PlacementGroup(pg_id, bundles, row.get("ready_oid"))
    raise ValueError(f"placement group {name_or_hex!r} not found")
