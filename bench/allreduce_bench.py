"""All-reduce sweep: hand-written peer-memory kernels vs NCCL.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29533 bench/allreduce_bench.py [--check] [--sizes 1024,...] [--json out.json]

Device-timed with CUDA events, max over ranks; bus bandwidth = 2 (N-1)/N * bytes / time.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def timed(fn, iters, warm=5):
    for _ in range(warm):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    t = torch.tensor([s.elapsed_time(e) / iters], device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item()) * 1e3     # microseconds


def stress(iters, rank, world, local):
    """Flag / epoch / parity protocols under skew: every rank delays each call by a different random time (a spin
    kernel on the stream), algorithms and sizes alternate, nothing synchronises with the host between calls.  Values
    are small integers, so every expected result is exact (also through the bf16 wire)."""
    import numpy as np
    from raydp_b200.comm import P2PAllToAll, P2PReducer
    dev = torch.device("cuda", local)
    rng = np.random.default_rng(1234 + rank)
    sizes = (1024, 48520, 1 << 20)
    algos = ("ll", "one_shot", "two_shot", "bf16")
    red = {(n, a): P2PReducer(torch.zeros(n, device=dev), dev, algo=a) for n in sizes for a in algos
           if not (a == "ll" and n * 4 > (1 << 20))}
    a2a = {c: P2PAllToAll(c, dev) for c in (256, 65536)}
    send = {c: torch.empty(world, c, device=dev) for c in a2a}
    dst_ids = torch.arange(world, device=dev, dtype=torch.float32).view(world, 1)
    keys = list(red)
    pending = []
    for it in range(iters):
        torch.cuda._sleep(int(rng.integers(0, 3_000_000)))            # up to ~1.5 ms of skew between the ranks
        if it % 5 == 4:
            c = (256, 65536)[(it // 5) % 2]
            send[c].copy_((rank * 1000.0 + dst_ids + (it % 8)).expand(world, c))
            got = a2a[c](send[c])
            want = torch.arange(world, device=dev, dtype=torch.float32).view(world, 1) * 1000.0 + rank + (it % 8)
            pending.append(("a2a", it, got.clone(), want.expand(world, c)))
        else:
            n, a = keys[it % len(keys)]
            r = red[(n, a)]
            r.grads.fill_(float(rank + 1 + (it % 8)))
            r.allreduce_()
            want = float(world * (world + 1) // 2 + world * (it % 8))
            pending.append((a, it, r.grads.clone(), want))
        if it % 20 == 19 or it == iters - 1:
            torch.cuda.synchronize()
            for name, i, got, want in pending:
                ok = torch.equal(got, want) if torch.is_tensor(want) else bool((got == want).all())
                assert ok, (name, i, rank, got.flatten()[:4].tolist(), want if not torch.is_tensor(want) else want.flatten()[:4].tolist())
            pending = []
    for r in red.values():
        r.close()
    for a in a2a.values():
        a.close()
    if rank == 0:
        print(f"STRESS OK {iters} calls", flush=True)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--sizes", default="256,4096,16384,48516,131072,262144,1048576,4194304,16777216,67108864,268435456")
    ap.add_argument("--json", default=None)
    ap.add_argument("--stress", type=int, default=0,
                    help="N back-to-back collectives (LL / one-shot / two-shot / bf16 all-reduce and the one-round "
                         "all-to-all, three sizes) with a random device-side delay on every rank before each call, "
                         "no host synchronisation in between; results checked exactly every 20 calls")
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    from raydp_b200.comm import NvlsReducer, P2PReducer, nvls_available, p2p_available
    assert p2p_available(), "peer access unavailable"
    use_nvls = nvls_available(torch.device("cuda", local))
    if rank == 0:
        print(f"# NVLS multicast support: {use_nvls}", flush=True)
    if args.stress:
        stress(args.stress, rank, world, local)
        dist.barrier()
        dist.destroy_process_group()
        return
    rows = []
    sampler = None
    if not args.check:
        from raydp_b200.utils.profiling import ClockSampler
        sampler = ClockSampler(local)
        sampler.start()
    for n in [int(x) for x in args.sizes.split(",")]:
        n = (n + 3) // 4 * 4
        g = torch.Generator(device="cuda").manual_seed(1000 + rank)
        x = torch.randn(n, device="cuda", generator=g)
        if args.check:
            ref = x.clone()
            dist.all_reduce(ref)
            for algo in ("one_shot", "two_shot", "ll"):
                if algo == "one_shot" and n > (1 << 22):
                    continue
                if algo == "ll" and n > (1 << 20):
                    continue
                red = P2PReducer(x.clone(), torch.device("cuda", local), algo=algo)
                for rep in range(3):
                    red.grads.copy_(x)
                    red.allreduce_()
                    torch.cuda.synchronize()
                    err = (red.grads - ref).abs().max().item()
                    assert err <= 1e-4 * max(1.0, ref.abs().max().item()), (algo, n, rep, err)
                red.close()
            red = P2PReducer(x.clone(), torch.device("cuda", local), algo="two_shot")
            for rep in range(3):
                red.next_send_buffer().copy_(x)
                got = red.allreduce_registered_()
                torch.cuda.synchronize()
                err = (got - ref).abs().max().item()
                assert err <= 1e-4 * max(1.0, ref.abs().max().item()), ("registered", n, rep, err)
            red.close()
            # bf16 on the wire: compare with the fp32 sum of the bf16-rounded inputs (what the kernels promise:
            # one rounding of the input, fp32 accumulation, one rounding of the result)
            n8 = n // 8 * 8
            xq = x[:n8].to(torch.bfloat16).float()
            refq = xq.clone()
            dist.all_reduce(refq)
            tol = 2.0 ** -8 * max(1.0, refq.abs().max().item())
            red = P2PReducer(x[:n8].clone(), torch.device("cuda", local), algo="bf16")
            for rep in range(3):
                red.grads.copy_(x[:n8])
                red.allreduce_()
                torch.cuda.synchronize()
                err = (red.grads - refq).abs().max().item()
                assert err <= tol, ("bf16", n, rep, err)
            red.close()
            if use_nvls:
                nv = NvlsReducer(n, torch.device("cuda", local))
                for rep in range(3):
                    nv.buffer.copy_(x)
                    nv.allreduce_()
                    torch.cuda.synchronize()
                    err = (nv.buffer - ref).abs().max().item()
                    assert err <= 1e-4 * max(1.0, ref.abs().max().item()), ("nvls", n, rep, err)
                nv.close()
                nv = NvlsReducer(n8, torch.device("cuda", local), wire="bf16")
                for rep in range(3):
                    nv.buffer.copy_(x[:n8])
                    nv.allreduce_()
                    torch.cuda.synchronize()
                    err = (nv.buffer - refq).abs().max().item()
                    assert err <= 2 * tol, ("nvls bf16", n, rep, err)     # the switch returns bf16: up to 1 ulp
                nv.close()
            continue
        iters = 200 if n <= (1 << 20) else (50 if n <= (1 << 24) else 10)
        buf = x.clone()
        t_nccl = timed(lambda: dist.all_reduce(buf), iters)
        res = {"floats": n, "bytes": n * 4, "nccl_us": t_nccl}
        if n >= (1 << 18):
            # bf16 on the wire (fp32 gradients in, fp32 result out): NCCL's side of it is cast -> all_reduce(bf16)
            # -> cast back, which is what DDP's bf16 compression hook does
            def nccl_bf16():
                h = buf.to(torch.bfloat16)
                dist.all_reduce(h)
                buf.copy_(h)
            res["nccl_bf16_us"] = timed(nccl_bf16, iters)
            n8 = n // 8 * 8
            red = P2PReducer(x[:n8].clone(), torch.device("cuda", local), algo="bf16")
            res["bf16_two_shot_us"] = timed(lambda: red.allreduce_(), iters)
            red.close()
            if use_nvls:
                nv = NvlsReducer(n8, torch.device("cuda", local), wire="bf16")
                nv.buffer.copy_(x[:n8])
                res["nvls_bf16_us"] = timed(lambda: nv.allreduce_(), iters)
                nv.close()
        for algo in ("one_shot", "two_shot", "ll"):
            if algo == "one_shot" and n > (1 << 22):
                continue
            if algo == "ll" and n > (1 << 20):
                continue
            red = P2PReducer(x.clone(), torch.device("cuda", local), algo=algo)
            res[f"{algo}_us"] = timed(lambda: red.allreduce_(), iters)
            red.close()
        if n >= (1 << 18):
            red = P2PReducer(x.clone(), torch.device("cuda", local), algo="two_shot")
            red.next_send_buffer().copy_(x)
            res["two_shot_registered_us"] = timed(lambda: red.allreduce_registered_(), iters)
            red.close()
        if use_nvls:
            nv = NvlsReducer(n, torch.device("cuda", local))
            nv.buffer.copy_(x)
            res["nvls_us"] = timed(lambda: nv.allreduce_(), iters)
            res["nvls_overlap_shape_us"] = timed(lambda: nv.allreduce_(overlap=True), iters)
            nv.close()
        best = min(v for k, v in res.items() if k.endswith("_us") and not k.startswith("nccl") and "bf16" not in k)
        wire16 = [v for k, v in res.items() if k.endswith("_us") and not k.startswith("nccl") and "bf16" in k]
        if wire16:
            res["bf16_best_us"] = min(wire16)
        factor = 2 * (world - 1) / world * n * 4
        res["p2p_best_us"] = best
        res["nccl_busbw_GBs"] = factor / t_nccl / 1e3
        res["p2p_busbw_GBs"] = factor / best / 1e3
        rows.append(res)
        if rank == 0:
            print(json.dumps(res), flush=True)
    if args.check and rank == 0:
        print("CHECK OK", flush=True)
    clocks = sampler.stop() if sampler is not None else None
    if rank == 0 and clocks is not None:
        print(json.dumps({"clocks": clocks}), flush=True)
    if args.json and rank == 0 and rows:
        with open(args.json, "w") as fh:
            json.dump({"world": world, "clocks": clocks, "rows": rows}, fh, indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
