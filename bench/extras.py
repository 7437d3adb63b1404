"""Short records for BASELINE.json configs 3-5, callable from ``bench.py`` inside one invocation:
DLRM (hybrid parallel vs replicated-embedding DDP), the 1 B-parameter MLP under ``FusedDataParallel``
vs stock DDP, and all-reduce / all-to-all points vs NCCL.  Every record carries its own clock sample
taken during its timed region; numbers are device-timed with CUDA events, max over ranks.

The full sweeps live in ``bench/{dlrm,bigmlp,allreduce,alltoall}_bench.py``; these are the same code paths
with few steps so the driver's end-of-round run records them next to the headline number.
"""
from __future__ import annotations

import gc
import time
from typing import Callable, Dict, List

import torch
import torch.distributed as dist
import torch.nn as nn

from raydp_b200.utils.profiling import ClockSampler


def _timed(step: Callable[[int], None], steps: int, warmup: int, world: int, local: int) -> Dict:
    sampler = ClockSampler(local)
    sampler.start()
    sampler.wait_first_sample()
    for i in range(warmup):
        step(i)
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()
    t_begin = time.time()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for i in range(steps):
        step(warmup + i)
    e.record()
    torch.cuda.synchronize()
    t_end = time.time()
    ms = torch.tensor([s.elapsed_time(e)], device="cuda", dtype=torch.float64)
    if world > 1:
        dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    return {"ms": float(ms.item()) / steps, "clocks": sampler.stop(t_begin, t_end)}


def _release():
    gc.collect()
    torch.cuda.empty_cache()


# ---------------------------------------------------------------------------------------------- DLRM
def dlrm_record(rank: int, world: int, local: int, steps: int = 20, warmup: int = 5, batch: int = 8192,
                rows: int = 100000) -> Dict:
    """BASELINE config 3: Criteo-shaped synthetic, 26 tables x `rows` x 32, per-GPU batch 8192."""
    from raydp_b200.models import DLRM
    from raydp_b200.parallel.hybrid import DistributedDLRM
    dev = torch.device("cuda", local)
    sizes = [rows] * 26
    torch.manual_seed(0)
    g = torch.Generator(device=dev).manual_seed(1 + rank)
    pool = []
    for _ in range(4):
        dense = torch.randn(batch, 13, device=dev, generator=g)
        idx = torch.randint(0, rows, (batch, 26), device=dev, generator=g, dtype=torch.int32)
        y = (torch.rand(batch, device=dev, generator=g) < 0.25).float()
        pool.append((dense, idx, y))
    out = {"metric": "DLRM train samples/sec", "per_gpu_batch": batch, "table_rows": rows, "n_gpus": world,
           "steps": steps, "warmup": warmup, "data": "synthetic, Criteo-shaped (13 dense + 26 categorical)"}
    # ours: table-wise sharded embeddings, lookup fused with the all-to-all, fused dense all-reduce + SGD
    model = DLRM(13, sizes, 32, sparse_grad=False)
    eng = DistributedDLRM(model.fused_spec(), batch, dev)
    eng.core.load_from_module(model)
    packed = []
    for dense, idx, y in pool:
        d = torch.zeros(batch, eng.core.kpad, device=dev, dtype=torch.bfloat16)
        d[:, :13] = dense.to(torch.bfloat16)
        packed.append((d, idx, y))
    r = _timed(lambda i: eng.train_step(*packed[i % 4], 0.01), steps, warmup, world, local)
    out["ours"] = {"ms_per_step": r["ms"], "value": world * batch / (r["ms"] / 1e3), "clocks": r["clocks"],
                   "parallelism": "dense dp + table-wise embedding mp"}
    eng.close()
    del eng, packed
    _release()
    # stand-in: replicated nn.Embedding tables under DDP, as the reference notebook does it
    model = model.to(dev)
    ddp = nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
    opt = torch.optim.SGD(ddp.parameters(), lr=0.01)
    crit = nn.BCEWithLogitsLoss()

    def step(i):
        dense, idx, y = pool[i % 4]
        with torch.autocast("cuda", dtype=torch.bfloat16):
            o = ddp(dense, idx.long())
        loss = crit(o[:, 0].float(), y)
        loss.item()
        opt.zero_grad()
        loss.backward()
        opt.step()
    r = _timed(step, steps, warmup, world, local)
    out["standin"] = {"ms_per_step": r["ms"], "value": world * batch / (r["ms"] / 1e3), "clocks": r["clocks"],
                      "parallelism": "ddp, replicated tables (stock torch + NCCL)"}
    out["ours_over_standin"] = out["ours"]["value"] / out["standin"]["value"]
    del ddp, model, opt, pool
    _release()
    return out


# -------------------------------------------------------------------------------------------- BigMLP
def bigmlp_record(rank: int, world: int, local: int, steps: int = 6, warmup: int = 3, width: int = 8192,
                  hidden: int = 16384, depth: int = 4, batch: int = 8192) -> Dict:
    """BASELINE config 4: ~1.07 B-parameter MLP, bucketed all-reduce fused with Adam under backward."""
    from raydp_b200.models import BigMLP
    from raydp_b200.parallel import FusedDataParallel
    dev = torch.device("cuda", local)
    torch.manual_seed(0)
    g = torch.Generator(device=dev).manual_seed(10 + rank)
    xs = [torch.randn(batch, width, device=dev, generator=g) for _ in range(2)]
    ys = [x[:, :4].sum(1, keepdim=True) for x in xs]
    crit = nn.MSELoss()
    out = {"metric": "BigMLP train step", "per_gpu_batch": batch, "n_gpus": world, "steps": steps, "warmup": warmup}
    with torch.device(dev):
        model = BigMLP(width, hidden, depth, fused=True)
    nparams = sum(p.numel() for p in model.parameters())
    flops = 6.0 * (nparams - width - 1) * batch
    dp = FusedDataParallel(model, torch.optim.Adam(model.parameters(), lr=1e-3), dev)

    def ours(i):
        loss = crit(dp(xs[i % 2].to(torch.bfloat16)), ys[i % 2])
        dp.backward_and_step(loss)
    r = _timed(ours, steps, warmup, world, local)
    out["params"] = nparams
    out["ours"] = {"ms_per_step": r["ms"], "tflops_per_gpu": flops / (r["ms"] / 1e3) / 1e12, "clocks": r["clocks"],
                   "grad_wire": "fp32"}
    dp.close()
    del dp, model
    _release()
    if world > 1:
        # same step with bf16 gradients on the wire (fp32 accumulate): the counterpart of DDP's bf16 compression hook
        with torch.device(dev):
            model = BigMLP(width, hidden, depth, fused=True)
        dp = FusedDataParallel(model, torch.optim.Adam(model.parameters(), lr=1e-3), dev, grad_wire="bf16")
        r = _timed(ours, steps, warmup, world, local)
        out["ours_bf16_wire"] = {"ms_per_step": r["ms"], "tflops_per_gpu": flops / (r["ms"] / 1e3) / 1e12,
                                 "clocks": r["clocks"], "grad_wire": "bf16"}
        dp.close()
        del dp, model
        _release()
    with torch.device(dev):
        model = BigMLP(width, hidden, depth, fused=False)
    ddp = nn.parallel.DistributedDataParallel(model, device_ids=[local]) if world > 1 else model
    opt = torch.optim.Adam(ddp.parameters(), lr=1e-3, fused=True)

    def standin(i):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            o = ddp(xs[i % 2])
        loss = crit(o.float(), ys[i % 2])
        opt.zero_grad()
        loss.backward()
        opt.step()
    r = _timed(standin, steps, warmup, world, local)
    out["standin"] = {"ms_per_step": r["ms"], "tflops_per_gpu": flops / (r["ms"] / 1e3) / 1e12, "clocks": r["clocks"],
                      "what": "stock torch: bf16 autocast, cuBLAS, DDP/NCCL, fused torch.optim.Adam"}
    out["ours_over_standin"] = out["standin"]["ms_per_step"] / out["ours"]["ms_per_step"]
    del ddp, model, opt, xs, ys
    _release()
    return out


# --------------------------------------------------------------------------------------- collectives
def allreduce_record(rank: int, world: int, local: int, sizes: List[int] = (16384, 49664, 131072, 262144, 4194304, 67108864)
                     ) -> Dict:
    """BASELINE config 5a, a few points (floats): 64 KB, 194 KB (the flagship's gradient), 512 KB, 1 MB,
    16 MB, 256 MB; ours = best of the peer-memory kernels (one-shot / two-shot / NVLS) vs ncclAllReduce."""
    from raydp_b200.comm import NvlsReducer, P2PReducer, nvls_available
    dev = torch.device("cuda", local)
    use_nvls = nvls_available(dev)
    rows = []
    for n in sizes:
        n = (n + 3) // 4 * 4
        x = torch.randn(n, device=dev)
        iters = 100 if n <= (1 << 20) else (20 if n <= (1 << 24) else 6)
        buf = x.clone()
        r = _timed(lambda i: dist.all_reduce(buf), iters, 5, world, local)
        row = {"floats": n, "bytes": n * 4, "nccl_us": r["ms"] * 1e3, "clocks": r["clocks"]}
        algos = ["one_shot", "two_shot"] if n <= (1 << 22) else ["two_shot"]
        if n <= (1 << 18):
            algos.append("ll")
        for algo in algos:
            try:
                red = P2PReducer(x.clone(), dev, algo=algo)
            except KeyError:
                continue
            row[f"{algo}_us"] = _timed(lambda i: red.allreduce_(), iters, 5, world, local)["ms"] * 1e3
            red.close()
        if n >= (1 << 22):
            # registered mode: the gradient is produced in the symmetric send buffer, no copy-in / copy-out
            # (what FusedDataParallel's buckets use below 4 ranks)
            red = P2PReducer(x.clone(), dev, algo="two_shot")
            red.next_send_buffer().copy_(x)
            row["two_shot_registered_us"] = _timed(lambda i: red.allreduce_registered_(), iters, 5, world, local)["ms"] * 1e3
            red.close()
        if use_nvls and n >= (1 << 16):
            nv = NvlsReducer(n, dev)
            nv.buffer.copy_(x)
            row["nvls_us"] = _timed(lambda i: nv.allreduce_(), iters, 5, world, local)["ms"] * 1e3
            nv.close()
        best = min(v for k, v in row.items() if k.endswith("_us") and k != "nccl_us")
        factor = 2 * (world - 1) / world * n * 4
        row.update(ours_best_us=best, nccl_busbw_GBs=factor / row["nccl_us"] / 1e3, ours_busbw_GBs=factor / best / 1e3)
        if n >= (1 << 22) and n % 8 == 0:
            # bf16 on the wire, fp32 gradients in and out (not part of ours_best_us, which is fp32 end to end)
            def nccl_bf16(i):
                h = buf.to(torch.bfloat16)
                dist.all_reduce(h)
                buf.copy_(h)
            wire = {"nccl_cast_allreduce_cast_us": _timed(nccl_bf16, iters, 5, world, local)["ms"] * 1e3}
            red = P2PReducer(x.clone(), dev, algo="bf16")
            wire["two_shot_us"] = _timed(lambda i: red.allreduce_(), iters, 5, world, local)["ms"] * 1e3
            red.close()
            if use_nvls:
                nv = NvlsReducer(n, dev, wire="bf16")
                nv.buffer.copy_(x)
                wire["nvls_us"] = _timed(lambda i: nv.allreduce_(), iters, 5, world, local)["ms"] * 1e3
                nv.close()
            row["bf16_wire"] = wire
        rows.append(row)
        del x, buf
        _release()
    return {"metric": "all-reduce fp32, ours (best peer-memory kernel) vs NCCL", "n_gpus": world, "rows": rows}


def alltoall_record(rank: int, world: int, local: int, sizes: List[int] = (16384, 4194304, 67108864)) -> Dict:
    """BASELINE config 5b, a few points (floats sent per rank): 64 KB, 16 MB, 256 MB."""
    from raydp_b200.comm import P2PAllToAll
    dev = torch.device("cuda", local)
    rows = []
    for total in sizes:
        per = max(4, (total // world + 3) // 4 * 4)
        x = torch.arange(world * per, device=dev, dtype=torch.float32) + 1e6 * rank
        ref = torch.empty_like(x)
        a2a = P2PAllToAll(per, dev)
        iters = 100 if total <= (1 << 22) else 10
        r = _timed(lambda i: dist.all_to_all_single(ref, x), iters, 5, world, local)
        t_ours = _timed(lambda i: a2a(x), iters, 5, world, local)["ms"] * 1e3
        a2a.close()
        nbytes = world * per * 4
        factor = (world - 1) / world * nbytes
        rows.append({"bytes": nbytes, "nccl_us": r["ms"] * 1e3, "ours_us": t_ours, "clocks": r["clocks"],
                     "nccl_busbw_GBs": factor / (r["ms"] * 1e3) / 1e3, "ours_busbw_GBs": factor / t_ours / 1e3})
        del x, ref
        _release()
    return {"metric": "all-to-all fp32, ours (peer stores) vs NCCL all_to_all_single", "n_gpus": world, "rows": rows}
