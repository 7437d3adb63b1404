"""DLRM step time of the hybrid engine on one GPU (or one rank per GPU under torchrun), per variant:
deterministic vs atomic reductions, bag sizes, table optimizers.  One JSON line per variant."""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=8192)
    ap.add_argument("--rows", type=int, default=100000)
    ap.add_argument("--steps", type=int, default=30)
    ap.add_argument("--warmup", type=int, default=6)
    ap.add_argument("--only", default=None, help="run only the variants whose name contains this text")
    args = ap.parse_args()
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        import torch.distributed as dist
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    from raydp_b200.models import DLRM
    from raydp_b200.parallel.hybrid import DistributedDLRM
    dev = torch.device("cuda", local)
    variants = [("atomic", dict(deterministic=False), None), ("deterministic", dict(deterministic=True), None),
                ("deterministic+adagrad", dict(deterministic=True, table_optimizer="adagrad"), None),
                ("deterministic+adam", dict(deterministic=True, table_optimizer="adam"), None),
                ("deterministic, bags of 4 (mean)", dict(deterministic=True), 4)]
    for name, kw, bag in variants:
        if args.only and args.only not in name:
            continue
        torch.manual_seed(0)
        sizes = [args.rows] * 26
        model = DLRM(13, sizes, 32, sparse_grad=False, bag_sizes=[bag] * 26 if bag else None,
                     pooling="mean" if bag else "sum")
        eng = DistributedDLRM(model.fused_spec(), args.batch, dev, **kw)
        eng.core.load_from_module(model)
        g = torch.Generator(device=dev).manual_seed(1 + local)
        pool = []
        for _ in range(4):
            d = torch.zeros(args.batch, eng.core.kpad, device=dev, dtype=torch.bfloat16)
            d[:, :13] = torch.randn(args.batch, 13, device=dev, generator=g).to(torch.bfloat16)
            idx = torch.randint(0, args.rows, (args.batch, 26 * (bag or 1)), device=dev, generator=g,
                                dtype=torch.int32)
            y = (torch.rand(args.batch, device=dev, generator=g) < 0.25).float()
            pool.append((d, idx, y))
        for i in range(args.warmup):
            eng.train_step(*pool[i % 4], 0.01)
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for i in range(args.steps):
            eng.train_step(*pool[i % 4], 0.01)
        b.record()
        torch.cuda.synchronize()
        ms = torch.tensor([a.elapsed_time(b) / args.steps], device=dev)
        if world > 1:
            dist.all_reduce(ms, op=dist.ReduceOp.MAX)
        if local == 0:
            print(json.dumps({"variant": name, "n_gpus": world, "per_gpu_batch": args.batch, "table_rows": args.rows,
                              "ms_per_step": round(ms.item(), 4),
                              "samples_per_s": round(world * args.batch / (ms.item() / 1e3))}), flush=True)
        eng.close()
        del eng, pool, model
        torch.cuda.empty_cache()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
