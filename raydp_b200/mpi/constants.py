"""Environment contract between the MPI driver and its rank processes.

Parity: reference ``python/raydp/mpi/constants.py:20-28``.
"""
MPI_TYPE = "raydp_mpi_type"
MPI_JOB_ID = "raydp_mpi_job_id"
MPI_DRIVER_HOST = "raydp_mpi_driver_host"
MPI_DRIVER_PORT = "raydp_mpi_driver_port"
# used by the built-in local launcher (no mpirun on PATH): it sets the rank variables itself
LOCAL_WORLD_RANK = "RAYDP_MPI_WORLD_RANK"
LOCAL_LOCAL_RANK = "RAYDP_MPI_LOCAL_RANK"

MAXIMUM_WAIT_TIME_OUT = 60
MAX_MESSAGE_BYTES = 50 * 1024 * 1024
