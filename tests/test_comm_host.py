"""Host-side logic of the communicator that needs no GPU: descriptor passing between rank processes
(unix sockets + SCM_RIGHTS, the transport of the cuMemCreate / cuMulticast handles) and the store-based
rendezvous with a deadline that replaces collective barriers in teardown (gloo, world_size 2)."""
import multiprocessing as mp
import os
import tempfile


def _rank_main(rank, world, port, tmpdir, q):
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank),
                       "WORLD_SIZE": str(world)})
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from raydp_b200 import comm
    try:
        # each rank offers two descriptors of files only it can name
        paths = [os.path.join(tmpdir, f"r{rank}_{k}.txt") for k in range(2)]
        for k, p in enumerate(paths):
            with open(p, "w") as fh:
                fh.write(f"rank {rank} file {k}")
        offered = [os.open(p, os.O_RDONLY) for p in paths]
        got = comm._exchange_fds(offered, rank, world, None, "t1")
        texts = {}
        for peer, fds in enumerate(got):
            texts[peer] = [os.pread(fd, 64, 0).decode() for fd in fds]
            if peer != rank:
                for fd in fds:
                    os.close(fd)
        ok_barrier = comm._host_barrier("unit_all", world, 20.0)
        # a rendezvous only one rank joins must give up at its deadline instead of hanging
        lonely = comm._host_barrier("unit_lonely", world, 0.5) if rank == 0 else None
        q.put((rank, texts, ok_barrier, lonely))
    finally:
        dist.barrier()
        dist.destroy_process_group()


def test_descriptor_exchange_and_bounded_rendezvous():
    from raydp_b200.utils import find_free_port
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = find_free_port()
    with tempfile.TemporaryDirectory() as tmpdir:
        procs = [ctx.Process(target=_rank_main, args=(r, 2, port, tmpdir, q)) for r in range(2)]
        for p in procs:
            p.start()
        results = {}
        for _ in range(2):
            rank, texts, ok_barrier, lonely = q.get(timeout=120)
            results[rank] = (texts, ok_barrier, lonely)
        for p in procs:
            p.join(60)
            assert p.exitcode == 0
    for rank in (0, 1):
        texts, ok_barrier, _ = results[rank]
        assert ok_barrier
        for peer in (0, 1):
            assert texts[peer] == [f"rank {peer} file 0", f"rank {peer} file 1"]
    assert results[0][2] is False
