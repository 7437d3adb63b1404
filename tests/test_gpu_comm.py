"""Peer-memory all-reduce kernels.

With one GPU the ranks are emulated inside one cooperative launch (blockIdx.y = rank): the same
device code (flag rounds, epoch parity, fused optimizer) runs, only the "peers" are local.  The
real multi-process / multi-GPU path is covered by ``test_multigpu_allreduce`` (needs >= 2 GPUs).
"""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("algo,n", [("one_shot", 48516), ("two_shot", 48516), ("two_shot", 3 * 1024 * 1024 + 8),
                                    ("ll", 48516), ("ll", 262144)])
def test_emulated_allreduce_matches_sum(world, algo, n):
    from raydp_b200.comm import LocalEmulation
    dev = torch.device("cuda")
    emu = LocalEmulation(world, n, dev)
    for step in range(3):            # several epochs: flag monotonicity and buffer parity
        g = torch.Generator(device=dev).manual_seed(100 * step + world)
        ins = [torch.randn(n, device=dev, generator=g) for _ in range(world)]
        outs = [torch.empty(n, device=dev) for _ in range(world)]
        ref = torch.stack(ins).sum(0) / world
        emu.allreduce(ins, outs, scale=1.0 / world, algo=algo)
        torch.cuda.synchronize()
        for o in outs:
            assert torch.allclose(o, ref, rtol=1e-5, atol=1e-5)


@pytest.mark.parametrize("world", [2, 8])
@pytest.mark.parametrize("n", [48520, 3 * 1024 * 1024 + 8])
def test_emulated_bf16_wire_allreduce_rounds_once_in_and_once_out(world, n):
    """bf16 on the wire, fp32 accumulate: the result is the bf16 rounding of the fp32 sum of the bf16-rounded
    inputs -- compared against exactly that, computed in fp32 by PyTorch."""
    from raydp_b200.comm import LocalEmulation
    dev = torch.device("cuda")
    emu = LocalEmulation(world, n, dev)
    for step in range(3):
        g = torch.Generator(device=dev).manual_seed(7 * step + world)
        ins = [torch.randn(n, device=dev, generator=g) for _ in range(world)]
        outs = [torch.empty(n, device=dev) for _ in range(world)]
        acc = torch.zeros(n, device=dev)
        for t in ins:                                 # rank order, like the kernel
            acc += t.to(torch.bfloat16).float()
        ref = (acc * (1.0 / world)).to(torch.bfloat16).float()
        emu.allreduce(ins, outs, scale=1.0 / world, algo="bf16")
        torch.cuda.synchronize()
        for o in outs:
            assert torch.equal(o, ref)
            assert (o - torch.stack(ins).mean(0)).abs().max().item() < 2e-2     # and close to the fp32 mean


def test_flag_protocol_stress_on_one_gpu():
    """Epoch-flag protocol under churn: 300 back-to-back all-reduces on emulated ranks, alternating
    one-shot / two-shot and message sizes without any host synchronisation in between (the parity
    double buffers and monotonically increasing epochs must keep successive calls apart)."""
    from raydp_b200.comm import LocalEmulation
    world, n_max = 4, 1 << 16
    dev = torch.device("cuda", 0)
    emu = {n: LocalEmulation(world, n, dev) for n in (1024, 8192, n_max)}
    torch.manual_seed(0)
    ins = {n: [torch.randn(n, device=dev) for _ in range(world)] for n in emu}
    outs = {n: [torch.empty(n, device=dev) for _ in range(world)] for n in emu}
    sizes = list(emu)
    for it in range(300):
        n = sizes[it % 3]
        for t in ins[n]:
            t.add_(1.0)                              # new data every call, enqueued on the same stream
        emu[n].allreduce(ins[n], outs[n], scale=1.0, algo=("one_shot", "two_shot", "ll", "ll")[it % 4])
        if it % 50 == 49:
            ref = torch.stack(ins[n]).sum(0)
            for o in outs[n]:
                assert torch.allclose(o, ref, rtol=1e-5, atol=1e-4), it
    torch.cuda.synchronize()


@pytest.mark.parametrize("algo", ["one_shot", "ll"])
def test_emulated_allreduce_fused_adam_matches_torch(algo):
    from raydp_b200.comm import LocalEmulation, make_opt_args
    dev = torch.device("cuda")
    world, n = 4, 48516
    emu = LocalEmulation(world, n, dev)
    torch.manual_seed(0)
    p0 = torch.randn(n, device=dev)
    params = [p0.clone() for _ in range(world)]
    ms = [torch.zeros(n, device=dev) for _ in range(world)]
    vs = [torch.zeros(n, device=dev) for _ in range(world)]
    step_dev = torch.zeros(1, device=dev)
    lr_dev = torch.full((1,), 1e-2, device=dev)
    p_ref = p0.clone().requires_grad_(True)
    opt = torch.optim.Adam([p_ref], lr=1e-2)
    for step in range(1, 4):
        grads = [torch.randn(n, device=dev) for _ in range(world)]
        p_ref.grad = torch.stack(grads).mean(0)
        opt.step()
        step_dev.fill_(float(step))
        opts = [make_opt_args(1, params[r], ms[r], vs[r], lr_dev=lr_dev, step_dev=step_dev) for r in range(world)]
        emu.allreduce(grads, grads, scale=1.0 / world, algo=algo, opts=opts)
        torch.cuda.synchronize()
    for r in range(world):
        assert torch.allclose(params[r], p_ref.detach(), rtol=1e-5, atol=1e-6)
        assert torch.equal(params[r], params[0])          # replicas stay bit-identical


@pytest.mark.parametrize("algo", ["one_shot", "ll"])
def test_wait_on_a_missing_peer_times_out_instead_of_hanging(algo):
    """Bounded peer waits: a 2-rank all-reduce whose "peer" never shows up (its region is plain local memory
    nobody writes) must end by itself after the deadline, leave a code in the host status word, make later
    waits return at once, and surface as CommTimeoutError on the host."""
    import ctypes
    import time
    from raydp_b200 import comm, ops
    dev = torch.device("cuda", 0)
    n = 4096
    comm.enable_timeouts(seconds=0.3)
    try:
        words = 2 * comm.LL_SOURCE_SLOTS * 2 * n if algo == "ll" else 4 * n + 64
        mine, ghost = torch.zeros(words, device=dev), torch.zeros(words, device=dev)
        x = torch.ones(n, device=dev)
        epoch = torch.zeros(1, dtype=torch.int32, device=dev)
        counters = torch.zeros(4, dtype=torch.int32, device=dev)
        t0 = time.time()
        launch = comm._launch_ll if algo == "ll" else comm._launch
        extra = () if algo == "ll" else (comm.ALGO["one_shot"],)
        launch(ops.lib(), [mine.data_ptr(), ghost.data_ptr()], n, 0, 2, [x], [x], [epoch], [counters], None, n, 1.0,
               *extra)
        torch.cuda.synchronize()
        assert 0.25 < time.time() - t0 < 10.0
        assert comm.comm_failed() != 0
        with pytest.raises(comm.CommTimeoutError):
            comm.check_status()
        t0 = time.time()                              # after a failure nothing waits any more
        launch(ops.lib(), [mine.data_ptr(), ghost.data_ptr()], n, 0, 2, [x], [x], [epoch], [counters], None, n, 1.0,
               *extra)
        torch.cuda.synchronize()
        assert time.time() - t0 < 0.25
    finally:
        comm._status_host.zero_()
        comm.enable_timeouts(seconds=60.0)            # re-arms and clears the device-side failure flag
    assert comm.comm_failed() == 0


@pytest.mark.multigpu
def test_multigpu_allreduce():
    n_gpus = torch.cuda.device_count()
    if n_gpus < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(n_gpus, 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29533", os.path.join(ROOT, "bench", "allreduce_bench.py"),
           "--check", "--sizes", "48516,1048576"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "CHECK OK" in out.stdout


@pytest.mark.multigpu
def test_multigpu_collectives_stress_with_random_delays():
    """the flag / epoch / parity protocols on real peers under skew -- 300 back-to-back calls
    (LL, one-shot, two-shot, bf16-wire all-reduce, one-round all-to-all; three sizes), every rank delayed by a
    different random spin before each call, no host synchronisation in between, exact results."""
    n_gpus = torch.cuda.device_count()
    if n_gpus < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(n_gpus, 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29537", os.path.join(ROOT, "bench", "allreduce_bench.py"),
           "--stress", "300"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "STRESS OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


@pytest.mark.multigpu
def test_multigpu_alltoall():
    n_gpus = torch.cuda.device_count()
    if n_gpus < 2:
        pytest.skip("needs >= 2 GPUs")
    world = min(n_gpus, 8)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}",
           "--master-addr", "127.0.0.1", "--master-port", "29534", os.path.join(ROOT, "bench", "alltoall_bench.py"),
           "--check", "--sizes", "1024,1048576"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout[-2000:] + out.stderr[-2000:]
    assert "CHECK OK" in out.stdout


def test_fused_data_parallel_single_rank_matches_torch():
    out = subprocess.run([sys.executable, os.path.join(ROOT, "bench", "bigmlp_bench.py"), "--check"],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "CHECK OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]


@pytest.mark.multigpu
def test_fused_data_parallel_multi_rank_matches_torch():
    n_gpus = torch.cuda.device_count()
    if n_gpus < 2:
        pytest.skip("needs >= 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29536", os.path.join(ROOT, "bench", "bigmlp_bench.py"), "--check"]
    out = subprocess.run(cmd, capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "CHECK OK" in out.stdout, out.stdout[-2000:] + out.stderr[-2000:]
