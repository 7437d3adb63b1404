"""MPI launcher: start / run / stop / restart, rank addresses, custom script + env, user PG.

Mirrors reference ``tests/test_mpi.py:28-129`` (there with real mpich on localhost; here the
built-in launcher stands in because the image has no ``mpirun``)."""
import pytest

from raydp_b200 import runtime as rt
from raydp_b200.mpi import MPIJobContext, WorkerContext, create_mpi_job


@pytest.mark.timeout(120)
def test_mpi_start_run_stop_restart(ray_cluster):
    job = create_mpi_job(job_name="test", world_size=2, num_cpus_per_process=1, num_processes_per_node=2,
                         timeout=5, mpi_type="mpich")
    job.start()

    def rank_fn(context: WorkerContext):
        return context.world_rank * 10 + context.local_rank

    assert job.run(rank_fn) == [0, 11]
    assert job.run(lambda ctx: ctx.job_id)[0] == job.job_id
    job.stop()
    # a stopped job can be started again (new ranks, function ids restart from 0)
    job.start()
    assert job.run(rank_fn) == [0, 11]
    job.stop()


@pytest.mark.timeout(120)
def test_mpi_context_manager_addresses_env_and_errors(ray_cluster):
    def prepare(context: MPIJobContext):
        context.add_env("RAYDP_TEST_FLAG", "hello")
        return ["mpirun", "-prepend-rank", "-hosts", ",".join(context.hosts), "-ppn", str(context.num_procs_per_node)]

    with create_mpi_job("ctx", 2, 1, 2, mpi_script_prepare_fn=prepare, timeout=5, mpi_type="intel_mpi") as job:
        addrs = job.get_rank_addresses()
        assert addrs == [rt.get_node_ip_address()] * 2                  # node IPs, like the reference
        ends = job.get_rank_endpoints()
        assert len(ends) == 2 and all(":" in a for a in ends) and ends[0] != ends[1]

        def read_env(ctx):
            import os
            return os.environ.get("RAYDP_TEST_FLAG"), ctx.node_ip
        out = job.run(read_env)
        assert [o[0] for o in out] == ["hello", "hello"]

        def uses_runtime(ctx):
            from raydp_b200 import runtime
            return runtime.get(runtime.put(ctx.world_rank + 100))
        assert job.run(uses_runtime) == [100, 101]

        def boom(ctx):
            raise ValueError("rank failure")
        with pytest.raises(Exception, match="function call failed"):
            job.run(boom)
    assert not job.started


@pytest.mark.timeout(120)
def test_mpi_with_user_placement_group(ray_cluster):
    before = rt.available_resources()["CPU"]
    pg = rt.placement_group([{"CPU": 2}], strategy="STRICT_PACK")
    assert pg.wait(10)
    with create_mpi_job("pg", 2, 1, 2, timeout=5, mpi_type="openmpi", placement_group=pg,
                        placement_group_bundle_indexes=[0]) as job:
        assert sorted(job.run(lambda ctx: ctx.world_rank)) == [0, 1]
    assert rt.placement_group_table(pg)["state"] == "CREATED"      # user-owned group is kept
    rt.remove_placement_group(pg)
    import time
    deadline = time.time() + 5
    while rt.available_resources()["CPU"] != before and time.time() < deadline:
        time.sleep(0.05)
    assert rt.available_resources()["CPU"] == before


def test_mpi_bad_arguments():
    with pytest.raises(AssertionError):
        create_mpi_job("bad", 3, 1, 2)
    with pytest.raises(Exception, match="not supported"):
        create_mpi_job("bad", 2, 1, 2, mpi_type="lam")
