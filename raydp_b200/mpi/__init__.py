"""MPI-on-runtime job launcher (reference ``python/raydp/mpi/__init__.py:25-94``, ``doc/mpi.md``)."""
from typing import Callable, List

from raydp_b200.mpi.mpi_job import (IntelMPIJob, MPICHJob, MPIJob, MPIJobContext, MPIType,  # noqa: F401
                                    OpenMPIJob)
from raydp_b200.mpi.mpi_worker import WorkerContext  # noqa: F401


def _get_mpi_type(mpi_type: str) -> MPIType:
    key = mpi_type.strip().lower()
    if key in ("openmpi", "open_mpi"):
        return MPIType.OPEN_MPI
    if key in ("intel_mpi", "intelmpi"):
        return MPIType.INTEL_MPI
    if key == "mpich":
        return MPIType.MPICH
    raise Exception(f"MPI type: {mpi_type} not supported now")


def create_mpi_job(job_name: str, world_size: int, num_cpus_per_process: int, num_processes_per_node: int,
                   mpi_script_prepare_fn: Callable = None, timeout: int = 1, mpi_type: str = "intel_mpi",
                   placement_group=None, placement_group_bundle_indexes: List[int] = None) -> MPIJob:
    """Create an MPI job: ``world_size`` ranks, ``num_processes_per_node`` per node, each reserving
    ``num_cpus_per_process`` CPUs through placeholder actors.  ``mpi_script_prepare_fn(context)`` may
    return a custom ``mpirun`` argv and add environment variables via ``context.add_env``."""
    kind = _get_mpi_type(mpi_type)
    cls = {MPIType.OPEN_MPI: OpenMPIJob, MPIType.INTEL_MPI: IntelMPIJob, MPIType.MPICH: MPICHJob}[kind]
    return cls(mpi_type=kind, job_name=job_name, world_size=world_size,
               num_cpus_per_process=num_cpus_per_process, num_processes_per_node=num_processes_per_node,
               mpi_script_prepare_fn=mpi_script_prepare_fn, timeout=timeout, placement_group=placement_group,
               placement_group_bundle_indexes=placement_group_bundle_indexes)


__all__ = ["create_mpi_job", "MPIJobContext", "WorkerContext", "MPIJob", "MPIType"]
