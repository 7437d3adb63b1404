"""DLRM throughput: hybrid-parallel fused path vs replicated-embedding DDP stand-in.

    python bench/dlrm_bench.py --batch 8192                                   # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29544 bench/dlrm_bench.py --batch 8192 [--impl standin]

Criteo-shaped synthetic data (13 dense + 26 categorical), random-init weights, per-GPU batch fixed
(weak scaling), device-timed with CUDA events, max over ranks.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402
import torch.nn as nn  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--rows", type=int, default=100000, help="rows per embedding table")
    ap.add_argument("--impl", default="ours", choices=["ours", "standin"])
    ap.add_argument("--breakdown", action="store_true",
                    help="after timing, print per-kernel device time of 4 steps (torch profiler; not a bench number)")
    ap.add_argument("--check", action="store_true",
                    help="ours only: graph-replayed steps vs eager steps from the same start, replicas bit-identical")
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    rank = dist.get_rank() if world > 1 else 0
    dev = torch.device("cuda", local)
    from raydp_b200.models import DLRM
    sizes = [args.rows] * 26
    torch.manual_seed(0)
    B = args.batch
    g = torch.Generator(device=dev).manual_seed(1 + rank)
    pool = []
    for _ in range(4):
        dense = torch.randn(B, 13, device=dev, generator=g)
        idx = torch.randint(0, args.rows, (B, 26), device=dev, generator=g, dtype=torch.int32)
        y = (torch.rand(B, device=dev, generator=g) < 0.25).float()
        pool.append((dense, idx, y))

    if args.check:
        from raydp_b200.parallel.hybrid import DistributedDLRM
        model = DLRM(13, [1000] * 26, 32, sparse_grad=False).to(dev)
        Bc = 1024
        gen = torch.Generator(device=dev).manual_seed(7 + rank)
        data = []
        for _ in range(8):
            d = torch.zeros(Bc, 16, device=dev, dtype=torch.bfloat16)
            d[:, :13] = torch.randn(Bc, 13, device=dev, generator=gen).to(torch.bfloat16)
            data.append((d, torch.randint(0, 1000, (Bc, 26), device=dev, generator=gen, dtype=torch.int32),
                         (torch.rand(Bc, device=dev, generator=gen) < 0.3).float()))
        finals = []
        for use_graph in (False, True):
            eng = DistributedDLRM(model.fused_spec(), Bc, dev, momentum=0.9, use_graph=use_graph)
            eng.core.load_from_module(model)
            for i, (d, idx, y) in enumerate(data):
                eng.train_step(d, idx, y, 0.05 if i < 5 else 0.02)
            torch.cuda.synchronize()
            params = eng.core.params.detach().clone()
            if world > 1:
                ref = params.clone()
                dist.broadcast(ref, src=0)
                assert torch.equal(ref, params), "dense replicas diverged"
            finals.append((params, [eng.core.tables[t].detach().clone() for t in eng.core.owned[:2]]))
            assert use_graph == (eng._graphs[0] is not None)
            eng.close()
        rel = ((finals[0][0] - finals[1][0]).norm() / finals[0][0].norm()).item()
        assert rel < 5e-3, rel
        # the step is deterministic (ordered reductions, fixed-point row combine, rank-ordered all-reduce):
        # replaying it from a graph must not change a single bit
        assert torch.equal(finals[0][0], finals[1][0]), "graph replay and eager launches differ in the dense parameters"
        for a, b in zip(finals[0][1], finals[1][1]):
            assert torch.equal(a, b), "graph replay and eager launches differ in the tables"
        for a, b in zip(finals[0][1], finals[1][1]):
            assert ((a - b).norm() / a.norm()).item() < 5e-3
        if rank == 0:
            print("CHECK OK", rel, flush=True)
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return

    if args.impl == "ours":
        from raydp_b200.parallel.hybrid import DistributedDLRM
        model = DLRM(13, sizes, 32, sparse_grad=False)
        eng = DistributedDLRM(model.fused_spec(), B, dev)
        eng.core.load_from_module(model.to(dev))
        del model
        packed = []
        for dense, idx, y in pool:
            d = torch.zeros(B, eng.core.kpad, device=dev, dtype=torch.bfloat16)
            d[:, :13] = dense.to(torch.bfloat16)
            packed.append((d, idx, y))

        def step(i):
            d, idx, y = packed[i % len(packed)]
            eng.train_step(d, idx, y, 0.01)
    else:
        model = DLRM(13, sizes, 32, sparse_grad=False).to(dev)     # replicated tables, dense grads (NCCL-reducible)
        ddp = nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
        opt = torch.optim.SGD(ddp.parameters(), lr=0.01)
        crit = nn.BCEWithLogitsLoss()

        def step(i):
            dense, idx, y = pool[i % len(pool)]
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = ddp(dense, idx.long())
            loss = crit(out[:, 0].float(), y)
            loss.item()
            opt.zero_grad()
            loss.backward()
            opt.step()

    for i in range(args.warmup):
        step(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(args.steps):
        step(i)
    e.record()
    torch.cuda.synchronize()
    ms = torch.tensor([s.elapsed_time(e)], device=dev)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    ms = float(ms.item())
    if rank == 0:
        print(json.dumps({"impl": args.impl, "metric": "DLRM train samples/sec", "n_gpus": world,
                          "per_gpu_batch": B, "table_rows": args.rows, "steps": args.steps,
                          "ms_per_step": ms / args.steps, "value": world * B * args.steps / (ms / 1e3),
                          "parallelism": "dense dp + table-wise embedding mp" if args.impl == "ours" else "ddp, replicated tables"}),
              flush=True)
    if args.breakdown:
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            for i in range(4):
                step(i)
            torch.cuda.synchronize()
        if rank == 0:
            rows = sorted(prof.key_averages(), key=lambda k: -k.device_time_total)
            total = sum(k.device_time_total for k in rows)
            print(f"  4 steps: {total / 1e3:.3f} ms of kernel time in {sum(k.count for k in rows)} kernels", flush=True)
            for k in rows[:22]:
                print(f"  {k.device_time_total / 4:8.1f} us/step  x{k.count // 4:<3d} {k.key[:88]}", flush=True)
    if args.impl == "ours":
        eng.close()
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
