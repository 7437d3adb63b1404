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


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--sizes", default="256,4096,16384,48516,131072,262144,1048576,4194304,16777216,67108864,268435456")
    ap.add_argument("--json", default=None)
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
    rows = []
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
            if use_nvls:
                nv = NvlsReducer(n, torch.device("cuda", local))
                for rep in range(3):
                    nv.buffer.copy_(x)
                    nv.allreduce_()
                    torch.cuda.synchronize()
                    err = (nv.buffer - ref).abs().max().item()
                    assert err <= 1e-4 * max(1.0, ref.abs().max().item()), ("nvls", n, rep, err)
                nv.close()
            continue
        iters = 200 if n <= (1 << 20) else (50 if n <= (1 << 24) else 10)
        buf = x.clone()
        t_nccl = timed(lambda: dist.all_reduce(buf), iters)
        res = {"floats": n, "bytes": n * 4, "nccl_us": t_nccl}
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
        best = min(v for k, v in res.items() if k.endswith("_us") and k != "nccl_us")
        factor = 2 * (world - 1) / world * n * 4
        res["p2p_best_us"] = best
        res["nccl_busbw_GBs"] = factor / t_nccl / 1e3
        res["p2p_busbw_GBs"] = factor / best / 1e3
        rows.append(res)
        if rank == 0:
            print(json.dumps(res), flush=True)
    if args.check and rank == 0:
        print("CHECK OK", flush=True)
    if args.json and rank == 0 and rows:
        with open(args.json, "w") as fh:
            json.dump({"world": world, "rows": rows}, fh, indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
