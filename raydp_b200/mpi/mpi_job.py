"""Driver side of the MPI-on-runtime launcher.

Parity: reference ``python/raydp/mpi/mpi_job.py:40-429`` -- ``MPIJob`` state machine
``start -> run* -> stop`` (restartable), the ``DriverService`` servicer (``:67-78``),
``MPIJobContext`` (``:90-115``), ``MPIWorkerPeer`` placeholder actors pinned by a STRICT_SPREAD
placement group to reserve CPUs and learn node IPs (``:47-49,192-222``), and per-flavour default
``mpirun`` command lines: OpenMPI ``mpirun --tag-output -H h1,h2 -N ppn`` (``:411-415``),
Intel MPI / MPICH ``mpirun -prepend-rank -hosts ... -ppn ...`` (``:418-429``).

When no ``mpirun`` binary is on PATH (this image) the job falls back to a built-in launcher that
starts the same rank processes itself with the same environment contract, so
``create_mpi_job(...).run(fn)`` keeps working; collectives inside ``fn`` are then the user's own
(e.g. ``torch.distributed``), exactly as with real MPI where RayDP never touches the data plane.
"""
from __future__ import annotations

import os
import shutil
import subprocess
import sys
import threading
from concurrent import futures
from enum import Enum, unique
from typing import Any, Callable, Dict, List, Optional

import cloudpickle
import grpc

from raydp_b200 import runtime
from raydp_b200.mpi import constants
from raydp_b200.mpi.mpi_worker import FailedFunc
from raydp_b200.mpi.network import (DRIVER_SERVICE, OPTIONS, create_insecure_channel, generic_handler,
                                    worker_stub)
from raydp_b200.mpi.utils import StoppableThread, run_cmd, terminate_process_group
from raydp_b200.runtime import protocol as _proto


@unique
class MPIType(Enum):
    OPEN_MPI = 0
    INTEL_MPI = 1
    MPICH = 2


class MPIWorkerPeer:
    """Placeholder actor: holds the CPUs of one node's ranks and reports the node address."""

    def __init__(self, job_id: str):
        self.job_id = job_id

    def get_node_ip(self) -> str:
        return runtime.get_node_ip_address()


class MPIJobContext:
    def __init__(self, job_id: str, hosts: List[str], num_procs_per_node: int):
        self.job_id = job_id
        self.hosts = hosts
        self.num_procs_per_node = num_procs_per_node
        self.env: Dict[str, str] = {}

    def add_env(self, key: str, value: str) -> None:
        self.env[key] = value

    def add_envs(self, envs: Dict[str, str]) -> None:
        self.env.update(envs)


class FunctionResults:
    def __init__(self, func_id: int, world_size: int):
        self.func_id = func_id
        self.remaining = world_size
        self.results: List[Any] = [None] * world_size
        self.done = threading.Event()


class MPIJob:
    def __init__(self, mpi_type: MPIType, job_name: str, world_size: int, num_cpus_per_process: int,
                 num_processes_per_node: int, mpi_script_prepare_fn: Optional[Callable] = None,
                 timeout: int = 1, placement_group=None, placement_group_bundle_indexes: Optional[List[int]] = None):
        assert world_size % num_processes_per_node == 0, (
            f"world_size: {world_size} should be multiple of num_processes_per_node: {num_processes_per_node}")
        self.mpi_type = mpi_type
        self.job_name = job_name
        self.world_size = world_size
        self.num_cpus_per_process = num_cpus_per_process
        self.num_processes_per_node = num_processes_per_node
        self.timeout = max(timeout, 1)
        self.script_prepare_fn = mpi_script_prepare_fn
        self.placement_group = placement_group
        self.placement_group_bundle_indexes = placement_group_bundle_indexes
        self.num_nodes = world_size // num_processes_per_node
        if placement_group_bundle_indexes is not None:
            assert len(placement_group_bundle_indexes) == self.num_nodes, (
                "placement_group_bundle_indexes should have world_size / num_processes_per_node entries")
        self._own_pg = None
        self.server = None
        self.peers: List[Any] = []
        self.node_addresses: List[str] = []
        self.workers: List[Any] = [None] * world_size
        self.rank_addresses: List[Optional[str]] = [None] * world_size
        self.mpirun_proc = None
        self.mpirun_check_thread: Optional[StoppableThread] = None
        self.local_procs: List[subprocess.Popen] = []
        self.lock = threading.RLock()
        self.register_cv = threading.Condition(self.lock)
        self.registered = 0
        self.func_id = 0
        self.func_result: Optional[FunctionResults] = None
        self.started = False
        self.failed = False
        self.job_id = ""

    # ------------------------------------------------------------------ DriverService
    def _register_worker(self, request):
        assert request["job_id"] == self.job_id, "job_id mismatch"
        with self.register_cv:
            self.registered += 1
            self.register_cv.notify_all()
        return {"node_addresses": self.node_addresses}

    def _register_worker_service(self, request):
        rank = request["world_rank"]
        address = f"{request['worker_ip']}:{request['worker_port']}"
        with self.register_cv:
            self.rank_addresses[rank] = address
            self.registered += 1
            self.register_cv.notify_all()
        env = {k: os.environ[k] for k in (_proto.ENV_HEAD, _proto.ENV_AUTH, _proto.ENV_NODE, _proto.ENV_NODE_IP,
                                         _proto.ENV_SESSION, _proto.ENV_NAMESPACE) if k in os.environ}
        if not env and runtime.is_initialized() and runtime._head is not None:
            head = runtime._head
            env = _proto.env_for_worker(head.address, head.authkey, head.head_node,
                                        runtime.get_node_ip_address(), head.session, f"mpi-{rank}", head.namespace)
            env = {k: v for k, v in env.items() if k.startswith("RAYDP_B200_")}
        return {"ray_address": runtime.session_info().get("address", ""), "redis_password": "",
                "runtime_env": env}

    def _register_func_result(self, request):
        with self.lock:
            fr = self.func_result
            if fr is None or fr.func_id != request["func_id"]:
                return {}
            fr.results[request["world_rank"]] = cloudpickle.loads(request["result"])
            fr.remaining -= 1
            if fr.remaining == 0:
                fr.done.set()
        return {}

    # ------------------------------------------------------------------ start
    def _start_network_service(self) -> None:
        self.server = grpc.server(futures.ThreadPoolExecutor(max_workers=4), options=OPTIONS)
        self.server.add_generic_rpc_handlers((generic_handler(DRIVER_SERVICE, {
            "RegisterWorker": self._register_worker,
            "RegisterWorkerService": self._register_worker_service,
            "RegisterFuncResult": self._register_func_result}),))
        self.driver_host = runtime.get_node_ip_address()
        self.driver_port = self.server.add_insecure_port(f"{self.driver_host}:0")
        self.server.start()

    def _start_peers(self) -> None:
        cpus = self.num_cpus_per_process * self.num_processes_per_node
        if self.placement_group is None:
            pg = runtime.placement_group([{"CPU": cpus}] * self.num_nodes,
                                         strategy="STRICT_SPREAD" if self.num_nodes > 1 else "PACK")
            runtime.get(pg.ready(), timeout=max(30, self.timeout))
            self._own_pg = pg
            indexes = list(range(self.num_nodes))
        else:
            pg = self.placement_group
            indexes = self.placement_group_bundle_indexes or list(range(self.num_nodes))
        Peer = runtime.remote(MPIWorkerPeer)
        self.peers = [Peer.options(num_cpus=cpus, placement_group=pg,
                                   placement_group_bundle_index=idx).remote(self.job_id) for idx in indexes]
        self.node_addresses = runtime.get([p.get_node_ip.remote() for p in self.peers])

    def get_default_mpirun_script(self, context: MPIJobContext) -> List[str]:
        raise NotImplementedError

    def _has_mpirun(self) -> bool:
        return shutil.which("mpirun") is not None and os.environ.get("RAYDP_MPI_FORCE_LOCAL") != "1"

    def _start_mpirun(self) -> None:
        self._start_peers()
        context = MPIJobContext(self.job_id, self.node_addresses, self.num_processes_per_node)
        env = dict(os.environ)
        env[constants.MPI_TYPE] = self.mpi_type.name
        env[constants.MPI_JOB_ID] = self.job_id
        env[constants.MPI_DRIVER_HOST] = self.driver_host
        env[constants.MPI_DRIVER_PORT] = str(self.driver_port)
        root = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
        env["PYTHONPATH"] = os.pathsep.join([root] + [p for p in sys.path if p and os.path.isdir(p)]
                                            + [env.get("PYTHONPATH", "")])
        worker_cmd = [sys.executable, "-u", "-m", "raydp_b200.mpi._rank_main"]
        if self._has_mpirun():
            script = (self.script_prepare_fn(context) if self.script_prepare_fn
                      else self.get_default_mpirun_script(context))
            env.update(context.env)
            cmd = " ".join(list(script) + worker_cmd)
            self.mpirun_proc, self.mpirun_check_thread = run_cmd(cmd, env=env, failed_callback=self._on_failed)
        else:
            if self.script_prepare_fn is not None:
                self.script_prepare_fn(context)         # users add env vars through the context
            env.update(context.env)
            for rank in range(self.world_size):
                renv = dict(env)
                renv[constants.LOCAL_WORLD_RANK] = str(rank)
                renv[constants.LOCAL_LOCAL_RANK] = str(rank % self.num_processes_per_node)
                proc, watchdog = run_cmd(worker_cmd, env=renv, failed_callback=self._on_failed, shell=False,
                                         prefix=f"[{rank}] ")
                self.local_procs.append(proc)
                if self.mpirun_check_thread is None:
                    self.mpirun_check_thread = watchdog
                else:
                    self._extra_watchdogs = getattr(self, "_extra_watchdogs", []) + [watchdog]
        self._wait_client_register(self.world_size)

    def _wait_client_register(self, target: int) -> None:
        timeout = max(self.timeout, constants.MAXIMUM_WAIT_TIME_OUT)
        with self.register_cv:
            ok = self.register_cv.wait_for(lambda: self.registered >= target or self.failed, timeout=timeout)
        if self.failed:
            raise Exception("mpirun exited with an error while the ranks were starting")
        if not ok:
            self._reset()
            raise Exception(f"Timeout to wait for MPI ranks to register ({self.registered}/{target})")

    def start(self) -> None:
        if self.started:
            return
        if not runtime.is_initialized():
            runtime.init()
        # the job id the ranks see IS the job name (reference mpi_job.py:260,281); a restart gets a new
        # driver port, so stale ranks of an earlier start cannot register with this one
        self.job_id = self.job_name
        self.failed = False
        self.registered = 0
        try:
            self._start_network_service()
            self._start_mpirun()                         # wait #1: every rank called RegisterWorker
            self._wait_client_register(2 * self.world_size)   # wait #2: every rank registered its service
            for rank, addr in enumerate(self.rank_addresses):
                self.workers[rank] = worker_stub(create_insecure_channel(addr))
            self.started = True
        except BaseException:
            self._reset()
            raise

    # ------------------------------------------------------------------ run / stop
    def run(self, mpi_func: Callable, timeout: Optional[float] = None) -> List[Any]:
        assert self.started, "the MPI job has not been started"
        with self.lock:
            func_id = self.func_id
            self.func_id += 1
            self.func_result = FunctionResults(func_id, self.world_size)
            fr = self.func_result
        blob = cloudpickle.dumps(mpi_func)
        for stub in self.workers:
            stub.RunFunction({"func_id": func_id, "func": blob})
        if not fr.done.wait(timeout):
            raise TimeoutError(f"MPI function {func_id} did not finish in {timeout}s")
        with self.lock:
            self.func_result = None
        if self.failed:
            raise Exception("function call failed")
        for rank, res in enumerate(fr.results):
            if isinstance(res, FailedFunc):
                raise Exception(f"function call failed on rank {rank}:\n{res.message}")
        return fr.results

    def get_rank_addresses(self) -> List[str]:
        """IP address of every rank's node, by rank (reference ``mpi_job.py:340-342``)."""
        assert self.started
        return [addr.rsplit(":", 1)[0] for addr in self.rank_addresses]

    def get_rank_endpoints(self) -> List[str]:
        """``ip:port`` of every rank's own RPC service, by rank."""
        assert self.started
        return list(self.rank_addresses)

    def _on_failed(self) -> None:
        self.failed = True
        with self.register_cv:
            self.register_cv.notify_all()
        fr = self.func_result
        if fr is not None:
            fr.done.set()

    def stop(self) -> None:
        for stub in self.workers:
            if stub is not None:
                try:
                    stub.Stop({}, timeout=2)
                except grpc.RpcError:
                    pass
        self._reset()

    def _reset(self) -> None:
        for wd in [self.mpirun_check_thread] + getattr(self, "_extra_watchdogs", []):
            if wd is not None:
                wd.stop()
        self.mpirun_check_thread = None
        self._extra_watchdogs = []
        if self.mpirun_proc is not None:
            terminate_process_group(self.mpirun_proc)
            self.mpirun_proc = None
        for proc in self.local_procs:
            try:
                proc.wait(timeout=2)
            except subprocess.TimeoutExpired:
                terminate_process_group(proc)
        self.local_procs = []
        if self.server is not None:
            self.server.stop(0.2)
            self.server = None
        for peer in self.peers:
            try:
                runtime.kill(peer)
            except Exception:  # noqa: BLE001
                pass
        self.peers = []
        if self._own_pg is not None:
            runtime.remove_placement_group(self._own_pg)
            self._own_pg = None
        self.workers = [None] * self.world_size
        self.rank_addresses = [None] * self.world_size
        self.registered = 0
        self.func_id = 0
        self.started = False

    def __enter__(self):
        self.start()
        return self

    def __exit__(self, exc_type, exc_val, exc_tb):
        self.stop()

    def __del__(self):
        try:
            if self.started:
                self.stop()
        except Exception:  # noqa: BLE001
            pass


class OpenMPIJob(MPIJob):
    def get_default_mpirun_script(self, context: MPIJobContext) -> List[str]:
        return ["mpirun", "--allow-run-as-root", "--tag-output", "-H", ",".join(context.hosts),
                "-N", f"{context.num_procs_per_node}"]


class IntelMPIJob(MPIJob):
    def get_default_mpirun_script(self, context: MPIJobContext) -> List[str]:
        return ["mpirun", "-prepend-rank", "-hosts", ",".join(context.hosts),
                "-ppn", f"{context.num_procs_per_node}"]


class MPICHJob(MPIJob):
    def get_default_mpirun_script(self, context: MPIJobContext) -> List[str]:
        return ["mpirun", "-prepend-rank", "-hosts", ",".join(context.hosts),
                "-ppn", f"{context.num_procs_per_node}"]
