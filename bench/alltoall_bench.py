"""All-to-all sweep: peer-store kernel vs NCCL ``all_to_all_single``.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29535 bench/alltoall_bench.py [--check] [--json out.json]

``bytes`` is the total each rank sends; bus bandwidth = (N-1)/N * bytes / time.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

from allreduce_bench import timed  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--check", action="store_true")
    ap.add_argument("--sizes", default="1024,16384,262144,4194304,67108864,268435456")
    ap.add_argument("--json", default=None)
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank, world = dist.get_rank(), dist.get_world_size()
    from raydp_b200.comm import P2PAllToAll
    rows = []
    sampler = None
    if not args.check:
        from raydp_b200.utils.profiling import ClockSampler
        sampler = ClockSampler(local)
        sampler.start()
    for total in [int(x) for x in args.sizes.split(",")]:
        per = max(4, (total // world + 3) // 4 * 4)
        x = torch.arange(world * per, device="cuda", dtype=torch.float32) + 1e6 * rank
        ref = torch.empty_like(x)
        a2a = P2PAllToAll(per, torch.device("cuda", local))
        if args.check:
            prev = None
            for it in range(5):                          # new data every call: the parity buffers must not mix
                x.add_(1.0)
                dist.all_to_all_single(ref, x)
                got = a2a(x)
                torch.cuda.synchronize()
                assert torch.equal(got.reshape(-1), ref), (total, rank, it)
                if prev is not None:                     # the previous result stays readable for one more call
                    assert torch.equal(prev[0].reshape(-1), prev[1]), (total, rank, it, "previous buffer clobbered")
                prev = (got, ref.clone())
            a2a.close()
            continue
        iters = 200 if total <= (1 << 22) else 20
        t_nccl = timed(lambda: dist.all_to_all_single(ref, x), iters)
        t_ours = timed(lambda: a2a(x), iters)
        a2a.close()
        nbytes = world * per * 4
        factor = (world - 1) / world * nbytes
        row = {"floats_total": world * per, "bytes": nbytes, "nccl_us": t_nccl, "p2p_us": t_ours,
               "nccl_busbw_GBs": factor / t_nccl / 1e3, "p2p_busbw_GBs": factor / t_ours / 1e3}
        rows.append(row)
        if rank == 0:
            print(json.dumps(row), flush=True)
    if args.check and rank == 0:
        print("CHECK OK", flush=True)
    clocks = sampler.stop() if sampler is not None else None
    if rank == 0 and clocks is not None:
        print(json.dumps({"clocks": clocks}), flush=True)
    if args.json and rank == 0 and rows:
        json.dump({"world": world, "clocks": clocks, "rows": rows}, open(args.json, "w"), indent=1)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
