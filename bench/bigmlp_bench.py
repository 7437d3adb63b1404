"""Large-MLP trainer benchmark: FusedDataParallel (tcgen05 FusedLinear + bucketed
peer-memory all-reduce fused with Adam, overlapped with backward) vs stock DDP + cuBLAS + torch.optim.

    python bench/bigmlp_bench.py --width 8192 --hidden 16384 --depth 4 --batch 8192
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29546 \
        bench/bigmlp_bench.py [--impl standin] [--no-overlap] [--check]

Reports samples/s, step time and -- for ours -- the exposed communication time (step time with the
all-reduce minus step time of the same step on one rank's worth of compute, i.e. ``--no-comm``).
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
    ap.add_argument("--width", type=int, default=8192)
    ap.add_argument("--hidden", type=int, default=16384)
    ap.add_argument("--depth", type=int, default=4)
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "standin"])
    ap.add_argument("--no-overlap", action="store_true")
    ap.add_argument("--bucket-mb", type=int, default=512)
    ap.add_argument("--check", action="store_true", help="small model: compare against fp32 torch + all_reduce")
    ap.add_argument("--breakdown", action="store_true",
                    help="after timing, print the per-kernel time of one step (torch profiler; not a bench number)")
    args = ap.parse_args()
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    rank = dist.get_rank() if world > 1 else 0
    from raydp_b200.models import BigMLP
    from raydp_b200.parallel import FusedDataParallel
    if args.check:
        args.width, args.hidden, args.depth, args.batch = 256, 512, 2, 1024
    torch.manual_seed(0)
    B = args.batch
    g = torch.Generator(device=dev).manual_seed(10 + rank)
    xs = [torch.randn(B, args.width, device=dev, generator=g) for _ in range(2)]
    ys = [x[:, :4].sum(1, keepdim=True) for x in xs]
    crit = nn.MSELoss()

    if args.impl == "ours":
        model = BigMLP(args.width, args.hidden, args.depth, fused=True).to(dev)
        nparams = sum(p.numel() for p in model.parameters())
        ref = None
        if args.check:
            ref = BigMLP(args.width, args.hidden, args.depth, fused=False).to(dev)
            sd = model.state_dict()
            ref.load_state_dict({k.replace("layers.", "layers."): v for k, v in _map_keys(sd, ref).items()})
            ref_opt = torch.optim.Adam(ref.parameters(), lr=1e-3)
        dp = FusedDataParallel(model, torch.optim.Adam(model.parameters(), lr=1e-3), dev, bucket_mb=args.bucket_mb,
                               overlap=not args.no_overlap)

        def step(i):
            x, y = xs[i % 2], ys[i % 2]
            loss = crit(dp(x.to(torch.bfloat16)), y)
            dp.backward_and_step(loss)
            return loss

        if args.check:
            for i in range(5):
                x, y = xs[i % 2], ys[i % 2]
                ref_opt.zero_grad()
                rl = crit(ref(x.to(torch.bfloat16).float()), y)
                rl.backward()
                if world > 1:
                    for p in ref.parameters():
                        dist.all_reduce(p.grad)
                        p.grad /= world
                ref_opt.step()
                ol = step(i)
                torch.cuda.synchronize()
                assert abs(ol.item() - rl.item()) < 0.05 * max(1.0, abs(rl.item())), (i, ol.item(), rl.item())
            w_ours = model.layers[0].weight.detach()
            w_ref = ref.layers[0].weight.detach()
            rel = ((w_ours - w_ref).norm() / w_ref.norm()).item()
            assert rel < 2e-2, rel
            if world > 1:
                chk = model.layers[0].weight.detach().clone()
                dist.broadcast(chk, src=0)
                assert torch.equal(chk, model.layers[0].weight.detach()), "replicas diverged"
            if rank == 0:
                print("CHECK OK", rel, flush=True)
            dp.close()
            _finish(world)
            return
    else:
        model = BigMLP(args.width, args.hidden, args.depth, fused=False).to(dev)
        nparams = sum(p.numel() for p in model.parameters())
        ddp = nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
        opt = torch.optim.Adam(ddp.parameters(), lr=1e-3)

        def step(i):
            x, y = xs[i % 2], ys[i % 2]
            with torch.autocast("cuda", dtype=torch.bfloat16):
                out = ddp(x)
            loss = crit(out.float(), y)
            opt.zero_grad()
            loss.backward()
            opt.step()
            return loss

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
    ms = float(ms.item()) / args.steps
    flops = 6.0 * (nparams - args.width - 1) * B
    if rank == 0:
        print(json.dumps({"impl": args.impl, "metric": "BigMLP train samples/sec", "n_gpus": world, "params": nparams,
                          "per_gpu_batch": B, "ms_per_step": ms, "value": world * B / (ms / 1e3),
                          "tflops_per_gpu": flops / (ms / 1e3) / 1e12, "overlap": not args.no_overlap}), flush=True)
    if args.breakdown:                      # every rank steps (collectives), rank 0 reports
        from torch.profiler import ProfilerActivity, profile
        with profile(activities=[ProfilerActivity.CUDA]) as prof:
            step(0)
            torch.cuda.synchronize()
        if rank == 0:
            if os.environ.get("RAYDP_B200_TRACE"):
                prof.export_chrome_trace(os.environ["RAYDP_B200_TRACE"])
            rows = sorted(prof.key_averages(), key=lambda k: -k.device_time_total)[:14]
            for k in rows:
                print(f"  {k.device_time_total / 1e3:8.2f} ms  x{k.count:<4d} {k.key[:90]}", flush=True)
    if args.impl == "ours":
        dp.close()
    _finish(world)


def _map_keys(sd, ref):
    """fused model: layers = [FL(relu), FL, FL(relu), FL ...]; stock: [Linear, ReLU, Linear, ...]."""
    out = {}
    for k, v in sd.items():
        if k.startswith("layers."):
            _, idx, name = k.split(".", 2)
            i = int(idx)
            j = (i // 2) * 3 + (0 if i % 2 == 0 else 2)
            out[f"layers.{j}.{name}"] = v
        else:
            out[k] = v
    return out


def _finish(world):
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
