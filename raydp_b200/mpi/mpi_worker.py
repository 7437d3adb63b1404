"""Rank process of an MPI job (run by ``mpirun`` -- or the built-in launcher -- on every rank).

Parity: reference ``python/raydp/mpi/mpi_worker.py:33-214``: read the job environment, find the
world / local rank (OpenMPI ``OMPI_COMM_WORLD_*``, Intel/MPICH ``PMI_RANK`` / ``MPI_LOCALRANKID``,
``:33-42``), register with the driver, start a ``WorkerService``, join the runtime, then execute
cloudpickled functions strictly in ``func_id`` order on a task-runner thread and post the results
back (``:63-96``).
"""
from __future__ import annotations

import os
import queue
import sys
import threading
import traceback
from concurrent import futures
from typing import Any, Callable

import cloudpickle
import grpc

from raydp_b200.mpi import constants
from raydp_b200.mpi.network import (OPTIONS, WORKER_SERVICE, create_insecure_channel, driver_stub,
                                    generic_handler)
from raydp_b200.mpi.utils import get_environ_value, get_node_ip_address


def get_rank(mpi_type: str):
    if os.environ.get(constants.LOCAL_WORLD_RANK) is not None:
        return int(os.environ[constants.LOCAL_WORLD_RANK]), int(os.environ[constants.LOCAL_LOCAL_RANK])
    if mpi_type == "OPEN_MPI":
        return int(os.environ["OMPI_COMM_WORLD_RANK"]), int(os.environ["OMPI_COMM_WORLD_LOCAL_RANK"])
    if mpi_type in ("INTEL_MPI", "MPICH"):
        return int(os.environ["PMI_RANK"]), int(os.environ.get("MPI_LOCALRANKID", os.environ["PMI_RANK"]))
    raise Exception(f"Not supported MPI type: {mpi_type}")


class WorkerContext:
    """What a user function receives on every rank."""

    def __init__(self, job_id: str, world_rank: int, local_rank: int, node_ip: str):
        self.job_id = job_id
        self.world_rank = world_rank
        self.local_rank = local_rank
        self.node_ip = node_ip


class FailedFunc:
    """Marker posted back when the user function raised (keeps the driver from hanging)."""

    def __init__(self, message: str):
        self.message = message


class TaskRunner(threading.Thread):
    def __init__(self, ctx: WorkerContext, post_result: Callable[[int, bytes], None]):
        super().__init__(name="raydp-mpi-task-runner", daemon=True)
        self.ctx = ctx
        self.post_result = post_result
        self.tasks: "queue.Queue" = queue.Queue()
        self.expected_id = 0
        self.should_stop = threading.Event()

    def submit(self, func_id: int, blob: bytes) -> None:
        self.tasks.put((func_id, blob))

    def run(self) -> None:
        while not self.should_stop.is_set():
            try:
                func_id, blob = self.tasks.get(timeout=0.1)
            except queue.Empty:
                continue
            if func_id != self.expected_id:        # a lost or duplicated request
                self.post_result(func_id, cloudpickle.dumps(FailedFunc(
                    f"rank {self.ctx.world_rank}: expected function id {self.expected_id}, got {func_id}")))
                continue
            self.expected_id += 1
            try:
                result = cloudpickle.loads(blob)(self.ctx)
            except BaseException:  # noqa: BLE001
                result = FailedFunc(traceback.format_exc())
            self.post_result(func_id, cloudpickle.dumps(result))


class MPIWorker:
    def __init__(self, mpi_type: str, job_id: str, driver_host: str, driver_port: int):
        self.mpi_type, self.job_id = mpi_type, job_id
        self.world_rank, self.local_rank = get_rank(mpi_type)
        self.driver = driver_stub(create_insecure_channel(f"{driver_host}:{driver_port}"))
        self.done = threading.Event()
        self.runner = None
        self.server = None

    def _post(self, func_id: int, blob: bytes) -> None:
        self.driver.RegisterFuncResult({"world_rank": self.world_rank, "func_id": func_id, "result": blob})

    def start(self) -> None:
        reply = self.driver.RegisterWorker({"job_id": self.job_id, "world_rank": self.world_rank})
        node_ip = get_node_ip_address(reply["node_addresses"])
        self.server = grpc.server(futures.ThreadPoolExecutor(max_workers=1), options=OPTIONS)
        self.server.add_generic_rpc_handlers((generic_handler(WORKER_SERVICE, {
            "RunFunction": self._run_function, "Stop": self._stop}),))
        port = self.server.add_insecure_port(f"{node_ip}:0")
        # the runner exists (and queues) before the driver can learn our address and call us
        self.runner = TaskRunner(WorkerContext(self.job_id, self.world_rank, self.local_rank, node_ip), self._post)
        self.server.start()
        reply = self.driver.RegisterWorkerService({"world_rank": self.world_rank, "worker_ip": node_ip,
                                                   "worker_port": port})
        # join the runtime the driver lives in (reference: ray.init(address=reply.ray_address))
        for key, value in (reply.get("runtime_env") or {}).items():
            os.environ[key] = value
        try:
            from raydp_b200 import runtime
            if reply.get("runtime_env"):
                runtime.init(address="auto")
        except Exception:  # noqa: BLE001 - the job still works without runtime access
            traceback.print_exc()
        self.runner.start()

    def _run_function(self, request):
        self.runner.submit(request["func_id"], request["func"])
        return {}

    def _stop(self, _request):
        self.done.set()
        return {}

    def wait_for_termination(self) -> None:
        self.done.wait()
        if self.runner is not None:
            self.runner.should_stop.set()
        if self.server is not None:
            self.server.stop(0.2)


def main() -> int:
    worker = MPIWorker(get_environ_value(constants.MPI_TYPE), get_environ_value(constants.MPI_JOB_ID),
                       get_environ_value(constants.MPI_DRIVER_HOST),
                       int(get_environ_value(constants.MPI_DRIVER_PORT)))
    worker.start()
    worker.wait_for_termination()
    return 0


if __name__ == "__main__":
    sys.exit(main())
