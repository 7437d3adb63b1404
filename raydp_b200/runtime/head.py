"""Control plane of the built-in actor runtime ("head").

It stands in for the parts of Ray core that the reference exercises:
named actors with resources and restart policy (reference ``RayExecutorUtils.java:46-61``:
``maxRestarts=-1, maxTaskRetries=-1, maxConcurrency=2``), tasks with lineage and retries
(reference ``spark/dataset.py:54`` ``max_retries=-1``), placement groups
(reference ``context.py:119-140``), an object directory over the shared-memory store, and
virtual nodes that can be added and removed for fault injection (reference
``test_spark_cluster.py:262-299``).  One head lives in the driver process as threads;
every actor / task worker is an OS process connected over a unix socket.
"""
from __future__ import annotations

import collections
import os
import shutil
import signal
import subprocess
import sys
import tempfile
import threading
import time
import traceback
from concurrent.futures import ThreadPoolExecutor
from multiprocessing.connection import Listener
from typing import Any, Dict, List, Optional, Tuple

import psutil

from raydp_b200 import store
from raydp_b200.runtime import protocol as P

PENDING, ALIVE, RESTARTING, DEAD = "PENDING_CREATION", "ALIVE", "RESTARTING", "DEAD"


class _Node:
    def __init__(self, node_id: str, ip: str, resources: Dict[str, float], is_head: bool = False):
        self.node_id = node_id
        self.ip = ip
        self.total = dict(resources)
        self.avail = dict(resources)
        self.alive = True
        self.is_head = is_head
        self.gpu_free: List[int] = list(range(int(resources.get("GPU", 0))))
        self.idle_workers: List["_Worker"] = []


class _Worker:
    def __init__(self, worker_id: str, node: _Node, proc: subprocess.Popen, actor_id: Optional[str]):
        self.worker_id = worker_id
        self.node = node
        self.proc = proc
        self.actor_id = actor_id
        self.conn = None
        self.send_lock = threading.Lock()
        self.ready = threading.Event()
        self.inflight: Dict[str, "_Call"] = {}
        self.dead = False

    def send(self, msg) -> bool:
        try:
            with self.send_lock:
                self.conn.send(msg)
            return True
        except (OSError, ValueError, AttributeError):
            return False


class _Object:
    __slots__ = ("oid", "event", "desc", "node_id", "lineage", "lost", "owner_actor")

    def __init__(self, oid: str):
        self.oid = oid
        self.event = threading.Event()
        self.desc: Optional[Tuple] = None
        self.node_id: Optional[str] = None
        self.lineage: Optional["_Task"] = None
        self.lost = False
        self.owner_actor: Optional[str] = None


class _Task:
    def __init__(self, oid, func_blob, args_blob, resources, max_retries, name, pg, bundle):
        self.oid = oid
        self.func_blob = func_blob
        self.args_blob = args_blob
        self.resources = resources
        self.max_retries = max_retries
        self.retries_left = max_retries
        self.name = name
        self.pg = pg
        self.bundle = bundle
        self.running = False


class _Call:
    def __init__(self, call_id, oid, method, args_blob):
        self.call_id = call_id
        self.oid = oid
        self.method = method
        self.args_blob = args_blob
        self.retries = 0


class _Actor:
    def __init__(self, actor_id, name, namespace, cls_blob, args_blob, class_name, opts):
        self.actor_id = actor_id
        self.name = name
        self.namespace = namespace
        self.cls_blob = cls_blob
        self.args_blob = args_blob
        self.class_name = class_name
        self.opts = opts
        self.resources: Dict[str, float] = opts["resources"]
        self.max_restarts = opts.get("max_restarts", 0)
        self.max_task_retries = opts.get("max_task_retries", 0)
        self.max_concurrency = opts.get("max_concurrency", 1)
        self.pg = opts.get("placement_group")
        self.bundle = opts.get("bundle_index", -1)
        self.state = PENDING
        self.num_restarts = 0
        self.worker: Optional[_Worker] = None
        self.queue: "collections.deque[_Call]" = collections.deque()
        self.lock = threading.RLock()
        self.node_id: Optional[str] = None
        self.gpus: List[int] = []
        self.death_cause = ""
        self.intentional_exit = False
        self.holds: Optional[Tuple[str, Any]] = None  # where resources were taken from
        self.create_oid = P.new_id()


class _PlacementGroup:
    def __init__(self, pg_id, bundles, strategy, name):
        self.pg_id = pg_id
        self.bundles = [dict(b) for b in bundles]
        self.strategy = strategy
        self.name = name
        self.state = "PENDING"
        self.bundle_nodes: List[Optional[str]] = [None] * len(bundles)
        self.bundle_avail: List[Dict[str, float]] = [dict(b) for b in bundles]
        self.ready_oid = P.new_id()


def _fits(avail: Dict[str, float], need: Dict[str, float]) -> bool:
    return all(avail.get(k, 0.0) + 1e-9 >= v for k, v in need.items() if v > 0)


def _take(avail: Dict[str, float], need: Dict[str, float]) -> None:
    for k, v in need.items():
        if v > 0:
            avail[k] = avail.get(k, 0.0) - v


def _give(avail: Dict[str, float], need: Dict[str, float]) -> None:
    for k, v in need.items():
        if v > 0:
            avail[k] = avail.get(k, 0.0) + v


class Head:
    def __init__(self, num_cpus: Optional[int] = None, num_gpus: Optional[int] = None,
                 resources: Optional[Dict[str, float]] = None, namespace: Optional[str] = None,
                 object_store_memory: Optional[int] = None, memory: Optional[int] = None,
                 log_to_driver: Optional[bool] = None):
        self.session = P.new_id()[:10]
        _reap_dead_sessions()
        self._marker = _session_marker(self.session, os.getpid())
        try:
            open(self._marker, "w").close()
        except OSError:
            self._marker = None
        self.namespace = namespace or f"ns-{self.session}"
        self.job_id = P.new_id()[:8]
        self.dir = tempfile.mkdtemp(prefix=f"raydp_b200_{self.session}_")
        self.log_dir = os.path.join(self.dir, "logs")
        os.makedirs(self.log_dir, exist_ok=True)
        self.address = os.path.join(self.dir, "head.sock")
        self.authkey = os.urandom(16)
        self._listener = Listener(self.address, family="AF_UNIX", authkey=self.authkey)
        try:  # accept() must wake up periodically so shutdown can end the accept thread
            self._listener._listener._socket.settimeout(0.5)
        except AttributeError:
            pass
        self._lock = threading.RLock()
        self._pool = ThreadPoolExecutor(max_workers=256, thread_name_prefix="rdp-head")
        self._nodes: Dict[str, _Node] = {}
        self._node_seq = 0
        self._actors: Dict[str, _Actor] = {}
        self._named: Dict[Tuple[str, str], str] = {}
        self._objects: Dict[str, _Object] = {}
        self._workers: Dict[str, _Worker] = {}
        self._pgs: Dict[str, _PlacementGroup] = {}
        self._task_queue: "collections.deque[_Task]" = collections.deque()
        self._pending_actors: "collections.deque[_Actor]" = collections.deque()
        self._shutdown = False
        self._freed: set = set()
        if num_cpus is None:
            num_cpus = os.cpu_count() or 1
        if num_gpus is None:
            num_gpus = _detect_gpus()
        head_res = {"CPU": float(num_cpus), "GPU": float(num_gpus),
                    "memory": float(memory or psutil.virtual_memory().total),
                    "object_store_memory": float(object_store_memory or 2 ** 34)}
        head_res.update(resources or {})
        self.head_node = self.add_node(head_res, is_head=True)
        self._accept_thread = threading.Thread(target=self._accept_loop, name="rdp-head-accept",
                                               daemon=True)
        self._accept_thread.start()
        # Worker output (the capability of the reference's Java agent, core/agent/.../Agent.java:49-97: stdout / stderr
        # of every worker go to a per-process log file that starts with a ":job_id:<id>" header).  With
        # log_to_driver the files are also followed and echoed on the driver's stderr, each line prefixed with the
        # worker it came from.
        if log_to_driver is None:
            log_to_driver = os.environ.get("RAYDP_B200_LOG_TO_DRIVER", "0") not in ("0", "", "false")
        self.log_to_driver = bool(log_to_driver)
        self._log_files: Dict[str, Tuple[str, str, int]] = {}      # path -> (tag, worker id, pid)
        self._log_offsets: Dict[str, int] = {}
        self._log_pass_lock = threading.Lock()
        self._log_thread = None
        if self.log_to_driver:
            self._log_thread = threading.Thread(target=self._log_monitor, name="rdp-head-logs", daemon=True)
            self._log_thread.start()

    # ------------------------------------------------------------------ nodes
    def add_node(self, resources: Dict[str, float], is_head: bool = False) -> str:
        with self._lock:
            self._node_seq += 1
            ip = f"127.0.0.{self._node_seq}"
            node_id = f"n{self._node_seq}"
            res = {k: float(v) for k, v in resources.items()}
            res.setdefault("CPU", 0.0)
            res.setdefault("memory", float(psutil.virtual_memory().total))
            res[f"node:{ip}"] = 1.0
            self._nodes[node_id] = _Node(node_id, ip, res, is_head)
        self._kick()
        return node_id

    def remove_node(self, node_id: str) -> None:
        """Kill a virtual node: its processes die and the objects it stored are lost."""
        with self._lock:
            node = self._nodes.get(node_id)
            if node is None or not node.alive:
                return
            node.alive = False
            victims = [w for w in self._workers.values() if w.node is node and not w.dead]
            for obj in self._objects.values():
                if obj.node_id == node_id and obj.desc is not None and obj.desc[0] == "shm":
                    obj.lost = True
            for pg in self._pgs.values():
                for i, nid in enumerate(pg.bundle_nodes):
                    if nid == node_id:
                        pg.bundle_nodes[i] = None
                        pg.bundle_avail[i] = dict(pg.bundles[i])
                        pg.state = "RESCHEDULING"
        for w in victims:
            _kill_proc(w.proc)
        store.delete_prefix(f"rdp_{self.session}_{node_id}_")
        self._kick()

    def nodes(self) -> List[dict]:
        with self._lock:
            return [{"NodeID": n.node_id, "NodeManagerAddress": n.ip, "Alive": n.alive,
                     "Resources": dict(n.total), "alive": n.alive} for n in self._nodes.values()]

    # ------------------------------------------------------------------ connections
    def _accept_loop(self) -> None:
        while not self._shutdown:
            try:
                conn = self._listener.accept()
            except TimeoutError:
                continue
            except (OSError, EOFError):
                if self._shutdown:
                    return
                continue
            except Exception:  # noqa: BLE001 - failed auth etc.
                continue
            threading.Thread(target=self._serve_conn, args=(conn,), daemon=True,
                             name="rdp-head-conn").start()

    def _serve_conn(self, conn) -> None:
        try:
            hello = conn.recv()
        except (EOFError, OSError):
            return
        if hello[0] == "worker":
            self._serve_worker(conn, hello[1])
            return
        info = {"node_id": hello[1], "worker_id": hello[2],
                "namespace": hello[3] if len(hello) > 3 else None}
        send_lock = threading.Lock()

        def run(req_id, op, kwargs):
            try:
                value = self.handle(op, kwargs, info)
                reply = (req_id, True, value)
            except BaseException as exc:  # noqa: BLE001
                try:
                    blob = P.dumps(exc)
                except Exception:  # noqa: BLE001
                    blob = P.dumps(RuntimeError(repr(exc)))
                reply = (req_id, False, blob)
            try:
                with send_lock:
                    conn.send(reply)
            except (OSError, ValueError):
                pass

        try:
            while True:
                req_id, op, kwargs = conn.recv()
                try:
                    self._pool.submit(run, req_id, op, kwargs)
                except RuntimeError:          # the head is shutting down: late requests are dropped
                    break
        except (EOFError, OSError):
            pass
        finally:
            try:
                conn.close()
            except OSError:
                pass

    def _serve_worker(self, conn, worker_id: str) -> None:
        with self._lock:
            worker = self._workers.get(worker_id)
        if worker is None:
            conn.close()
            return
        worker.conn = conn
        worker.ready.set()
        try:
            while True:
                msg = conn.recv()
                if msg[0] == "done":
                    _, call_id, desc, node_id = msg
                    self._on_call_done(worker, call_id, desc, node_id)
                elif msg[0] == "exit_actor":
                    with self._lock:
                        actor = self._actors.get(worker.actor_id or "")
                    if actor is not None:
                        actor.intentional_exit = True
        except (EOFError, OSError):
            pass
        finally:
            self._on_worker_death(worker)

    # ------------------------------------------------------------------ dispatch
    def handle(self, op: str, kw: dict, info: dict) -> Any:
        return getattr(self, "op_" + op)(info, **kw)

    # objects ----------------------------------------------------------------
    def _obj(self, oid: str) -> _Object:
        with self._lock:
            obj = self._objects.get(oid)
            if obj is None:
                obj = self._objects[oid] = _Object(oid)
            return obj

    def _complete(self, oid: str, desc: Tuple, node_id: Optional[str]) -> None:
        obj = self._obj(oid)
        obj.desc = desc
        obj.node_id = node_id
        obj.lost = False
        obj.event.set()

    def op_put(self, info, oid: str, desc: Tuple, node_id: str, owner_actor=None):
        obj = self._obj(oid)
        obj.owner_actor = owner_actor
        self._complete(oid, desc, node_id)
        return True

    def op_get(self, info, oid: str, timeout: Optional[float] = None):
        if oid in self._freed:
            raise P.ObjectLostError(f"object {oid} was freed (its owner cleaned it up)")
        obj = self._obj(oid)
        deadline = None if timeout is None else time.time() + timeout
        while True:
            if obj.lost:
                self._reconstruct(obj)
            remaining = None if deadline is None else max(0.0, deadline - time.time())
            if not obj.event.wait(remaining if remaining is not None else 1.0):
                if deadline is not None and time.time() >= deadline:
                    raise P.GetTimeoutError(f"get timed out after {timeout}s for object {oid}")
                if self._shutdown:
                    raise P.RuntimeErrorBase("runtime is shut down")
                continue
            if obj.lost:
                continue
            return obj.desc

    def op_report_lost(self, info, oid: str):
        obj = self._obj(oid)
        with self._lock:
            if obj.desc is not None and obj.desc[0] == "shm" and not store.exists(obj.desc[1]):
                obj.lost = True
        return True

    def _reconstruct(self, obj: _Object) -> None:
        with self._lock:
            if not obj.lost:
                return
            task = obj.lineage
            if task is None:
                obj.desc = P.error_descriptor(P.ObjectLostError(
                    f"object {obj.oid} was lost and has no lineage to rebuild it"))
                obj.lost = False
                obj.event.set()
                return
            if task.running:
                obj.lost = False
                return
            if task.max_retries >= 0 and task.retries_left <= 0:
                obj.desc = P.error_descriptor(P.ObjectLostError(
                    f"object {obj.oid} was lost and its task has no retries left"))
                obj.lost = False
                obj.event.set()
                return
            if task.max_retries >= 0:
                task.retries_left -= 1
            obj.lost = False
            obj.event.clear()
            obj.desc = None
            task.running = True
            self._task_queue.append(task)
        self._kick()

    def op_wait(self, info, oids: List[str], num_returns: int, timeout: Optional[float]):
        deadline = None if timeout is None else time.time() + timeout
        while True:
            ready = [o for o in oids if self._obj(o).event.is_set() and not self._obj(o).lost]
            if len(ready) >= num_returns or (deadline is not None and time.time() >= deadline):
                ready = ready[:num_returns]
                return ready, [o for o in oids if o not in ready]
            time.sleep(0.002)

    def op_free(self, info, oids: List[str]):
        with self._lock:
            objs = [self._objects.pop(o, None) for o in oids]
            self._freed.update(oids)
        for obj in objs:
            if obj is not None and obj.desc is not None and obj.desc[0] == "shm":
                store.delete(obj.desc[1])
        return True

    def op_object_locations(self, info, oids: List[str]):
        out = []
        with self._lock:
            for o in oids:
                obj = self._objects.get(o)
                node = self._nodes.get(obj.node_id) if obj is not None and obj.node_id else None
                out.append(node.ip if node is not None else None)
        return out

    # tasks ------------------------------------------------------------------
    def op_submit_task(self, info, func_blob, args_blob, resources, max_retries, name,
                       placement_group=None, bundle_index=-1, prefer_ip=None):
        oid = P.new_id()
        task = _Task(oid, func_blob, args_blob, resources, max_retries, name,
                     placement_group, bundle_index)
        task.prefer_ip = prefer_ip
        obj = self._obj(oid)
        obj.lineage = task
        task.running = True
        with self._lock:
            self._task_queue.append(task)
        self._kick()
        return oid

    def _kick(self) -> None:
        if not self._shutdown:
            self._pool.submit(self._schedule)

    def _schedule(self) -> None:
        """Place queued actors and tasks wherever resources allow."""
        try:
            self._place_pgs()
            with self._lock:
                actors = list(self._pending_actors)
                self._pending_actors.clear()
            for actor in actors:
                if not self._try_start_actor(actor):
                    with self._lock:
                        self._pending_actors.append(actor)
            with self._lock:
                tasks = list(self._task_queue)
                self._task_queue.clear()
            leftover = []
            for task in tasks:
                if not self._try_start_task(task):
                    leftover.append(task)
            if leftover:
                with self._lock:
                    self._task_queue.extendleft(reversed(leftover))
        except Exception:  # noqa: BLE001
            traceback.print_exc()

    def _reserve(self, need: Dict[str, float], pg_id: Optional[str], bundle: int,
                 prefer: Optional[str] = None):
        """Take ``need`` from a placement-group bundle or from a node; returns (node, holds)."""
        with self._lock:
            if pg_id:
                pg = self._pgs.get(pg_id)
                if pg is None or pg.state != "CREATED":
                    return None
                idxs = [bundle] if bundle is not None and bundle >= 0 else range(len(pg.bundles))
                for i in idxs:
                    node = self._nodes.get(pg.bundle_nodes[i] or "")
                    if node is None or not node.alive:
                        continue
                    if _fits(pg.bundle_avail[i], need):
                        _take(pg.bundle_avail[i], need)
                        return node, ("pg", (pg_id, i))
                return None
            order = [n for n in self._nodes.values() if n.alive]
            if prefer:
                order.sort(key=lambda n: n.node_id != prefer)
            for node in order:
                if _fits(node.avail, need):
                    _take(node.avail, need)
                    return node, ("node", node.node_id)
            return None

    def _release(self, need: Dict[str, float], holds) -> None:
        if holds is None:
            return
        with self._lock:
            kind, where = holds
            if kind == "pg":
                pg = self._pgs.get(where[0])
                if pg is not None:
                    _give(pg.bundle_avail[where[1]], need)
            else:
                node = self._nodes.get(where)
                if node is not None:
                    _give(node.avail, need)
        self._kick()

    def _spawn_worker(self, node: _Node, actor: Optional[_Actor], gpus: List[int]) -> _Worker:
        worker_id = P.new_id()[:12]
        env = P.env_for_worker(self.address, self.authkey, node.node_id, node.ip, self.session,
                               worker_id, actor.namespace if actor else self.namespace)
        if actor is not None:
            env["RAYDP_B200_ACTOR_ID"] = actor.actor_id
        if gpus:
            # Assigned device ids are advertised, not enforced through CUDA_VISIBLE_DEVICES: the
            # peer-memory collectives map every rank's GPU, so all devices must stay visible.
            env["RAYDP_B200_GPUS"] = ",".join(str(g) for g in gpus)
        env.setdefault("PYTHONUNBUFFERED", "1")
        if "OMP_NUM_THREADS" not in os.environ:
            # One intra-op thread per reserved CPU (1 for tasks and zero-CPU actors): several
            # workers share a node, and unbounded OpenMP pools oversubscribe it.
            cpus = actor.resources.get("CPU", 0) if actor is not None else 1
            env["OMP_NUM_THREADS"] = str(max(1, int(cpus)))
        extra_env =(actor.opts.get("env") if actor else None) or {}
        env.update({k: str(v) for k, v in extra_env.items()})
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        # workers see the driver's import path (functions/classes are pickled by reference)
        driver_path = [p for p in sys.path if p and os.path.isdir(p)]
        env["PYTHONPATH"] = os.pathsep.join([root] + driver_path + [env.get("PYTHONPATH", "")])
        tag = f"actor-{actor.name or actor.actor_id[:8]}" if actor else "task"
        log_path = os.path.join(self.log_dir, f"worker-{tag}-{worker_id}.log")
        log = open(log_path, "ab")
        header = f":job_id:{self.job_id}\n" + (f":actor_name:{actor.name or actor.class_name}\n" if actor else ":task_worker:\n")
        log.write(header.encode())
        log.flush()
        proc = subprocess.Popen([sys.executable, "-m", "raydp_b200.runtime.worker"],
                                env=env, stdin=subprocess.DEVNULL,
                                stdout=None if os.environ.get("RAYDP_B200_WORKER_STDOUT") else log,
                                stderr=subprocess.STDOUT if not os.environ.get(
                                    "RAYDP_B200_WORKER_STDOUT") else None,
                                start_new_session=True)
        log.close()
        worker = _Worker(worker_id, node, proc, actor.actor_id if actor else None)
        worker.log_path = log_path
        with self._lock:
            self._workers[worker_id] = worker
            self._log_files[log_path] = (tag, worker_id, proc.pid)
        return worker

    # ------------------------------------------------------------------ worker logs
    def _log_pass(self, final: bool = False) -> bool:
        """Echo what the worker log files gained since the last pass as ``(tag pid=N) line`` on the driver's stderr
        (whole lines; ``final`` also flushes an unterminated last line).  False when stderr is gone."""
        with self._lock:
            files = dict(self._log_files)
        offsets = self._log_offsets
        for path, (tag, _wid, pid) in files.items():
            try:
                with open(path, "rb") as fh:
                    fh.seek(offsets.get(path, 0))
                    chunk = fh.read()
            except OSError:
                continue
            cut = len(chunk) if final else chunk.rfind(b"\n") + 1
            offsets[path] = offsets.get(path, 0) + cut
            for line in chunk[:cut].decode("utf-8", "replace").splitlines():
                if line.startswith((":job_id:", ":actor_name:", ":task_worker:")):
                    continue
                try:
                    sys.stderr.write(f"({tag} pid={pid}) {line}\n")
                except (OSError, ValueError):
                    return False
        return True

    def _log_monitor(self) -> None:
        """Follow every worker log file until shutdown (which runs the last pass itself, before the files go)."""
        while not self._shutdown:
            with self._log_pass_lock:
                if self._shutdown or not self._log_pass():
                    return
            time.sleep(0.2)

    def op_worker_logs(self, info, name=None, tail_bytes=16384):
        """``{log file: {"tag", "worker_id", "pid", "job_id", "text"}}`` for workers whose tag contains ``name``."""
        out = {}
        with self._lock:
            files = dict(self._log_files)
        for path, (tag, wid, pid) in files.items():
            if name and name not in tag:
                continue
            try:
                with open(path, "rb") as fh:
                    fh.seek(0, os.SEEK_END)
                    size = fh.tell()
                    fh.seek(max(0, size - int(tail_bytes)))
                    text = fh.read().decode("utf-8", "replace")
            except OSError:
                continue
            out[path] = {"tag": tag, "worker_id": wid, "pid": pid, "job_id": self.job_id, "text": text}
        return out

    def _try_start_task(self, task: _Task) -> bool:
        prefer = None
        if getattr(task, "prefer_ip", None):
            with self._lock:
                prefer = next((n.node_id for n in self._nodes.values()
                               if n.alive and n.ip == task.prefer_ip), None)
        got = self._reserve(task.resources, task.pg, task.bundle, prefer=prefer)
        if got is None:
            return False
        node, holds = got
        with self._lock:
            worker = None
            while node.idle_workers:
                cand = node.idle_workers.pop()
                if not cand.dead:
                    worker = cand
                    break
        if worker is None:
            worker = self._spawn_worker(node, None, [])
        call = _Call(P.new_id(), task.oid, "__task__", None)
        call.task = task  # type: ignore[attr-defined]
        call.holds = holds  # type: ignore[attr-defined]

        def launch():
            if not self._wait_ready(worker):
                self._release(task.resources, holds)
                self._task_failed(task, "task worker failed to start")
                return
            with self._lock:
                worker.inflight[call.call_id] = call
            if not worker.send(("exec", call.call_id, "task", task.func_blob, task.args_blob, task.oid)):
                pass  # death handler requeues
        self._pool.submit(launch)
        return True

    def _task_failed(self, task: _Task, why: str) -> None:
        """A worker died under the task: retry (lineage) or fail the object."""
        with self._lock:
            if task.max_retries < 0 or task.retries_left > 0:
                if task.max_retries >= 0:
                    task.retries_left -= 1
                self._task_queue.append(task)
                retry = True
            else:
                retry = False
                task.running = False
        if retry:
            time.sleep(0.05)
            self._kick()
        else:
            self._complete(task.oid, P.error_descriptor(P.TaskError(f"task {task.name} failed: {why}")), None)

    # actors -----------------------------------------------------------------
    def op_create_actor(self, info, cls_blob, args_blob, class_name, opts):
        name = opts.get("name")
        namespace = opts.get("namespace") or info.get("namespace") or self.namespace
        actor_id = P.new_id()
        actor = _Actor(actor_id, name, namespace, cls_blob, args_blob, class_name, opts)
        with self._lock:
            if name:
                key = (namespace, name)
                old = self._named.get(key)
                if old and self._actors[old].state != DEAD:
                    if opts.get("get_if_exists"):
                        return old, self._actors[old].create_oid
                    raise ValueError(f"actor name {name!r} is already taken in namespace {namespace!r}")
                self._named[key] = actor_id
            self._actors[actor_id] = actor
            self._pending_actors.append(actor)
        self._kick()
        return actor_id, actor.create_oid

    def _wait_ready(self, worker: _Worker, timeout: float = 60.0) -> bool:
        """Wait for a spawned worker's handshake, giving up at once when the head shuts down."""
        deadline = time.time() + timeout
        while not worker.ready.wait(0.25):
            if self._shutdown or worker.dead or time.time() > deadline:
                return False
        return not (worker.dead or self._shutdown)

    def _try_start_actor(self, actor: _Actor) -> bool:
        if actor.state == DEAD or self._shutdown:
            return True
        got = self._reserve(actor.resources, actor.pg, actor.bundle, prefer=None)
        if got is None:
            return False
        node, holds = got
        ngpu = int(actor.resources.get("GPU", 0))
        with self._lock:
            short = ngpu > len(node.gpu_free)
            gpus = [] if short else [node.gpu_free.pop(0) for _ in range(ngpu)]
        if short:                                   # ids of a dying actor are not back yet: try again later
            self._release(actor.resources, holds)
            return False
        with self._lock:
            actor.gpus = gpus
            actor.holds = holds
            actor.node_id = node.node_id
        worker = self._spawn_worker(node, actor, gpus)
        actor.worker = worker

        def launch():
            if not self._wait_ready(worker):
                return  # death handler takes over
            ctx = {"actor_id": actor.actor_id, "name": actor.name, "namespace": actor.namespace,
                   "was_restarted": actor.num_restarts > 0, "max_concurrency": actor.max_concurrency,
                   "job_id": self.job_id, "gpus": gpus}
            call = _Call(P.new_id(), actor.create_oid, "__init__", None)
            with self._lock:
                worker.inflight[call.call_id] = call
            worker.send(("exec", call.call_id, "create", actor.cls_blob, actor.args_blob, ctx))
        self._pool.submit(launch)
        return True

    def _on_call_done(self, worker: _Worker, call_id: str, desc: Tuple, node_id: str) -> None:
        with self._lock:
            call = worker.inflight.pop(call_id, None)
        if call is None:
            return
        task = getattr(call, "task", None)
        if task is not None:
            task.running = False
            self._release(task.resources, call.holds)
            with self._lock:
                if not worker.dead and worker.node.alive:
                    worker.node.idle_workers.append(worker)
            self._complete(call.oid, desc, node_id)
            self._kick()
            return
        actor = self._actors.get(worker.actor_id or "")
        if call.method == "__init__" and actor is not None:
            if desc[0] == "error":
                actor.death_cause = "constructor raised"
                actor.intentional_exit = True
                actor.max_restarts = 0
                self._complete(call.oid, desc, node_id)
                _kill_proc(worker.proc)
                return
            with actor.lock:
                actor.state = ALIVE
                queued = list(actor.queue)
                actor.queue.clear()
                for c in queued:
                    self._send_call(actor, c)
            self._complete(call.oid, desc, node_id)
            return
        self._complete(call.oid, desc, node_id)

    def _send_call(self, actor: _Actor, call: _Call) -> None:
        worker = actor.worker
        with self._lock:
            worker.inflight[call.call_id] = call
        worker.send(("exec", call.call_id, "call", call.method, call.args_blob, call.oid))

    def op_call_actor(self, info, actor_id, method, args_blob):
        actor = self._actors.get(actor_id)
        oid = P.new_id()
        self._obj(oid)
        if actor is None or actor.state == DEAD:
            why = actor.death_cause if actor else "unknown actor"
            self._complete(oid, P.error_descriptor(P.ActorDiedError(
                f"actor {actor_id} is dead ({why})")), None)
            return oid
        call = _Call(P.new_id(), oid, method, args_blob)
        with actor.lock:
            if actor.state == ALIVE:
                self._send_call(actor, call)
            else:
                actor.queue.append(call)
        return oid

    def _on_worker_death(self, worker: _Worker) -> None:
        with self._lock:
            if worker.dead:
                return
            worker.dead = True
            self._workers.pop(worker.worker_id, None)
            if worker in worker.node.idle_workers:
                worker.node.idle_workers.remove(worker)
            inflight = list(worker.inflight.values())
            worker.inflight.clear()
        try:
            worker.proc.wait(timeout=5)
        except Exception:  # noqa: BLE001
            _kill_proc(worker.proc)
        if self._shutdown:
            return
        if worker.actor_id is None:
            for call in inflight:
                task = getattr(call, "task", None)
                if task is not None:
                    self._release(task.resources, call.holds)
                    self._task_failed(task, "worker process died")
            return
        actor = self._actors.get(worker.actor_id)
        if actor is None or actor.worker is not worker:
            return
        # device ids go back BEFORE the GPU resource count does: a new actor that wins the freed resource
        # must also find a free device id (else two ranks ended up on one GPU after a group restart)
        with self._lock:
            node = worker.node
            node.gpu_free.extend(actor.gpus)
            node.gpu_free.sort()
            actor.gpus = []
        self._release(actor.resources, actor.holds)
        with self._lock:
            actor.holds = None
        can_restart = (not actor.intentional_exit and
                       (actor.max_restarts < 0 or actor.num_restarts < actor.max_restarts))
        with actor.lock:
            if can_restart:
                actor.state = RESTARTING
                actor.num_restarts += 1
                retry, fail = [], []
                for call in inflight:
                    if call.method == "__init__":
                        continue
                    if actor.max_task_retries < 0 or call.retries < actor.max_task_retries:
                        call.retries += 1
                        retry.append(call)
                    else:
                        fail.append(call)
                actor.queue.extendleft(reversed(retry))
                actor.create_oid = actor.create_oid  # same ref stays valid
            else:
                actor.state = DEAD
                if not actor.death_cause:
                    actor.death_cause = "exit_actor" if actor.intentional_exit else "process died"
                fail = [c for c in inflight] + list(actor.queue)
                actor.queue.clear()
        for call in fail:
            self._complete(call.oid, P.error_descriptor(P.ActorDiedError(
                f"actor {actor.name or actor.actor_id} died ({actor.death_cause or 'process died'})")), None)
        if can_restart:
            with self._lock:
                self._pending_actors.append(actor)
            self._kick()
        else:
            with self._lock:
                if actor.name and self._named.get((actor.namespace, actor.name)) == actor.actor_id:
                    del self._named[(actor.namespace, actor.name)]
            self._kick()

    def op_get_actor(self, info, name, namespace=None):
        namespace = namespace or info.get("namespace") or self.namespace
        with self._lock:
            actor_id = self._named.get((namespace, name))
            if actor_id is None:
                # fall back to a unique match in any namespace (single-job runtime)
                matches = [aid for (ns, nm), aid in self._named.items() if nm == name]
                actor_id = matches[0] if len(matches) == 1 else None
            if actor_id is None or self._actors[actor_id].state == DEAD:
                raise ValueError(f"Failed to look up actor with name '{name}'")
            actor = self._actors[actor_id]
            return actor_id, actor.class_name, actor.create_oid

    def op_kill_actor(self, info, actor_id, no_restart=True):
        actor = self._actors.get(actor_id)
        if actor is None:
            return False
        if no_restart:
            actor.intentional_exit = True
            actor.death_cause = "killed"
        with self._lock:
            if actor in self._pending_actors:
                self._pending_actors.remove(actor)
        worker = actor.worker
        if worker is not None and not worker.dead:
            _kill_proc(worker.proc)
        elif no_restart:
            actor.state = DEAD
        return True

    def op_actor_state(self, info, actor_id):
        actor = self._actors.get(actor_id)
        return None if actor is None else actor.state

    def op_list_actors(self, info, filters=None):
        out = []
        with self._lock:
            for a in self._actors.values():
                rec = {"actor_id": a.actor_id, "name": a.name or "", "state": a.state,
                       "class_name": a.class_name, "namespace": a.namespace,
                       "required_resources": {k: v for k, v in a.resources.items() if v > 0},
                       "node_id": a.node_id, "num_restarts": a.num_restarts,
                       "pid": a.worker.proc.pid if a.worker is not None else None,
                       "death_cause": a.death_cause,
                       "node_ip": self._nodes[a.node_id].ip if a.node_id in self._nodes else None}
                out.append(rec)
        for key, op_, val in (filters or []):
            if op_ == "=":
                out = [r for r in out if r.get(key) == val]
            elif op_ == "!=":
                out = [r for r in out if r.get(key) != val]
        return out

    # placement groups -------------------------------------------------------
    def op_create_pg(self, info, bundles, strategy, name):
        if strategy not in ("PACK", "SPREAD", "STRICT_PACK", "STRICT_SPREAD"):
            raise ValueError(f"Invalid placement group strategy {strategy}")
        if not bundles or any(not b or any(v < 0 for v in b.values()) for b in bundles):
            raise ValueError("Bundles must be a non-empty list of non-empty resource dicts")
        pg = _PlacementGroup(P.new_id(), bundles, strategy, name)
        self._obj(pg.ready_oid)
        with self._lock:
            self._pgs[pg.pg_id] = pg
        self._place_pgs()
        return pg.pg_id, pg.ready_oid

    def _place_pgs(self) -> None:
        with self._lock:
            for pg in self._pgs.values():
                if pg.state in ("CREATED", "REMOVED"):
                    continue
                if self._place_one_pg(pg):
                    pg.state = "CREATED"
                    self._complete(pg.ready_oid, ("inline", P.dumps(True)), None)

    def _place_one_pg(self, pg: _PlacementGroup) -> bool:
        todo = [i for i, n in enumerate(pg.bundle_nodes) if n is None]
        if not todo:
            return True
        nodes = [n for n in self._nodes.values() if n.alive]
        trial = {n.node_id: dict(n.avail) for n in nodes}
        used_nodes = {nid for nid in pg.bundle_nodes if nid is not None}
        plan: Dict[int, str] = {}

        def place(i, candidates) -> bool:
            for n in candidates:
                if _fits(trial[n.node_id], pg.bundles[i]):
                    _take(trial[n.node_id], pg.bundles[i])
                    plan[i] = n.node_id
                    used_nodes.add(n.node_id)
                    return True
            return False

        if pg.strategy == "STRICT_PACK":
            ok = False
            for n in nodes:
                trial_n = dict(n.avail)
                if all(_fits_take(trial_n, pg.bundles[i]) for i in todo):
                    for i in todo:
                        plan[i] = n.node_id
                    ok = True
                    break
            if not ok:
                return False
        elif pg.strategy == "PACK":
            for i in todo:
                cands = sorted(nodes, key=lambda n: (n.node_id not in used_nodes, n.node_id))
                if not place(i, cands):
                    return False
        elif pg.strategy == "STRICT_SPREAD":
            for i in todo:
                cands = [n for n in nodes if n.node_id not in used_nodes]
                if not place(i, cands):
                    return False
        else:  # SPREAD: distinct nodes when possible
            for k, i in enumerate(todo):
                fresh = [n for n in nodes if n.node_id not in used_nodes]
                rest = [n for n in nodes if n.node_id in used_nodes]
                rot = k % max(1, len(rest))
                if not place(i, fresh + rest[rot:] + rest[:rot]):
                    return False
        for i, nid in plan.items():
            _take(self._nodes[nid].avail, pg.bundles[i])
            pg.bundle_nodes[i] = nid
            pg.bundle_avail[i] = dict(pg.bundles[i])
        return True

    def op_remove_pg(self, info, pg_id):
        with self._lock:
            pg = self._pgs.get(pg_id)
            if pg is None or pg.state == "REMOVED":
                return False
            for i, nid in enumerate(pg.bundle_nodes):
                node = self._nodes.get(nid or "")
                if node is not None:
                    _give(node.avail, pg.bundles[i])
                pg.bundle_nodes[i] = None
            pg.state = "REMOVED"
            victims = [a for a in self._actors.values() if a.pg == pg_id and a.state != DEAD]
        for a in victims:
            self.op_kill_actor(info, a.actor_id, True)
        self._kick()
        return True

    def op_pg_table(self, info, pg_id=None):
        with self._lock:
            def row(pg):
                return {"placement_group_id": pg.pg_id, "name": pg.name or "", "state": pg.state,
                        "strategy": pg.strategy, "ready_oid": pg.ready_oid,
                        "bundles": {i: dict(b) for i, b in enumerate(pg.bundles)},
                        "bundles_to_node_id": {i: n for i, n in enumerate(pg.bundle_nodes)}}
            if pg_id is not None:
                pg = self._pgs.get(pg_id)
                return row(pg) if pg else {}
            return {pg.pg_id: row(pg) for pg in self._pgs.values()}

    # cluster info -----------------------------------------------------------
    def op_resources(self, info, which="available"):
        out: Dict[str, float] = collections.defaultdict(float)
        with self._lock:
            for n in self._nodes.values():
                if not n.alive:
                    continue
                src = n.avail if which == "available" else n.total
                for k, v in src.items():
                    if which == "available" and v <= 1e-9:
                        continue
                    out[k] += v
        return dict(out)

    def op_nodes(self, info):
        return self.nodes()

    def op_add_node(self, info, resources):
        return self.add_node(resources)

    def op_remove_node(self, info, node_id):
        self.remove_node(node_id)
        return True

    def op_session_info(self, info):
        return {"session": self.session, "namespace": self.namespace, "dir": self.dir,
                "log_dir": self.log_dir, "job_id": self.job_id, "address": self.address,
                "head_node_ip": self._nodes[self.head_node].ip}

    def op_node_ip(self, info, node_id):
        node = self._nodes.get(node_id)
        return node.ip if node else "127.0.0.1"

    def op_ping(self, info):
        return "pong"

    # shutdown ---------------------------------------------------------------
    def shutdown(self) -> None:
        if self._shutdown:
            return
        self._shutdown = True
        with self._lock:
            workers = list(self._workers.values())
        for w in workers:
            w.send(("exit",))
        deadline = time.time() + 1.0
        for w in workers:
            try:
                w.proc.wait(timeout=max(0.0, deadline - time.time()))
            except Exception:  # noqa: BLE001
                pass
        for w in workers:
            _kill_proc(w.proc)
            w.dead = True
            w.ready.set()           # releases launch threads still waiting for the worker's handshake
        try:
            self._listener.close()
        except OSError:
            pass
        for obj in list(self._objects.values()):
            obj.event.set()
        store.delete_prefix(f"rdp_{self.session}_")
        if self._marker:
            try:
                os.unlink(self._marker)
            except OSError:
                pass
        self._pool.shutdown(wait=False, cancel_futures=True)
        if self.log_to_driver:
            with self._log_pass_lock:            # what the workers printed in their last moments
                self._log_pass(final=True)
        shutil.rmtree(self.dir, ignore_errors=True)


_SHM_DIR = "/dev/shm"


def _session_marker(session: str, pid: int) -> str:
    # not an object of the store (different prefix), so delete_prefix("rdp_<session>_") leaves it alone
    return os.path.join(_SHM_DIR, f"rdphead_{session}_{pid}")


def _reap_dead_sessions() -> int:
    """Objects are shared-memory segments that outlive their creator.  A driver that is killed (SIGKILL,
    ``os._exit``, a test timeout) never runs ``shutdown`` and would leave its whole session behind in /dev/shm.
    Every head leaves a marker carrying its pid; a new head removes the segments of sessions whose head is
    gone.  Returns the number of segments removed."""
    removed = 0
    try:
        names = os.listdir(_SHM_DIR)
    except OSError:
        return 0
    for name in names:
        if not name.startswith("rdphead_"):
            continue
        try:
            _, session, pid = name.split("_")
            pid = int(pid)
        except ValueError:
            continue
        alive = True
        try:
            os.kill(pid, 0)
        except ProcessLookupError:
            alive = False
        except PermissionError:
            alive = True
        if alive:
            continue
        removed += max(0, store.delete_prefix(f"rdp_{session}_"))
        try:
            os.unlink(os.path.join(_SHM_DIR, name))
        except OSError:
            pass
    return removed


def _fits_take(avail, need) -> bool:
    if _fits(avail, need):
        _take(avail, need)
        return True
    return False


def _kill_proc(proc: subprocess.Popen) -> None:
    if proc.poll() is not None:
        return
    try:
        os.killpg(proc.pid, signal.SIGKILL)
    except (ProcessLookupError, PermissionError, OSError):
        try:
            proc.kill()
        except OSError:
            pass


def _detect_gpus() -> int:
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis is not None and vis.strip() != "":
        return len([v for v in vis.split(",") if v.strip() and v.strip() != "-1"])
    try:
        return len([d for d in os.listdir("/proc/driver/nvidia/gpus")])
    except OSError:
        return 0
