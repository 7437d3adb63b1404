#!/usr/bin/env python
"""Headline benchmark: TorchEstimator training throughput (samples/s) of the NYC-taxi MLP.

Metric and config come from BASELINE.json: "TorchEstimator train samples/sec @1/2/4/8 B200
(NYC-taxi MLP ...)", config 2 (bf16 activations, data parallel, synthetic rows of the NYC-taxi
shape: 20 float features + fare label, random-init weights).

    python bench.py --gpus 1 --steps 50 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 \
        --master-port 29511 bench.py --gpus 8 --steps 50 --warmup 5

The model is the class a user of the reference writes -- ``NYC_Model`` of the reference's
``examples/pytorch_nyctaxi.py:69-92``, a plain ``nn.Module`` with no knowledge of this framework
(``tests/user_models.py``); the engine is picked by ``TorchEstimator`` itself (``engine`` in the record).

Numbers in the one JSON line (weak scaling: the per-GPU batch is fixed):

* ``value``       device-timed (CUDA events, max over ranks) samples/s of the full training step
                  (forward, backward, gradient all-reduce, optimizer) on device-resident synthetic
                  batches; the working set of one step (>150 MB of activations) exceeds the 126 MB
                  L2 and inputs rotate over a pool, so no step re-reads a warm cache.
* ``e2e.value``   the same metric through the public API, ``TorchEstimator.fit`` on a Dataset of Arrow
                  blocks in the shared-memory object store: every step DMA's its inputs from pinned
                  host memory, runs the ingest kernel, trains, and copies the running loss/metric
                  sums back to pinned host memory.
* ``standin``     the reference-equivalent loop on stock parts, run in the SAME invocation with the same
                  N / steps / warm-up and its own clock samples: ``device`` (bf16 autocast, cuBLAS,
                  DDP/NCCL, fused ``torch.optim.Adam`` on device-resident batches; eager and, if it
                  compiles, ``torch.compile``) and ``e2e`` (CPU batch assembly, blocking H2D, per-step
                  ``loss.item()``: the reference's loop, ``torch/estimator.py:272-293``).
                  ``vs_baseline`` = ``value`` / best stand-in device number; ``vs_baseline_detail`` has
                  both ratios.  These are stand-ins, never "the reference measured" (it cannot be
                  installed here).
* ``replica_check`` (N > 1) parameters bit-identical across ranks after the steps, and max-abs difference
                  against the same steps reduced with NCCL instead of the peer-memory kernel.
* ``e2e_actor``   one run of the default user path: ``init_spark`` -> ``spark.range``/``withColumn`` frame
                  -> ``fit_on_spark`` with TrainWorker ACTORS (not SPMD), conversion time reported apart.
* ``extra``       short records of BASELINE configs 3-5 (DLRM, 1 B-parameter MLP, all-reduce and
                  all-to-all points vs NCCL), each with its own clocks (``bench/extras.py``).

``--impl reference`` reports that the reference cannot be installed in this image;
``--impl standin`` prints the stand-in e2e loop as its own line.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time
import traceback

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))       # user_models.py: the reference's NYC_Model class

FEATURES = 20
METRIC = "TorchEstimator train samples/sec (NYC-taxi MLP)"


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference", "standin"])
    ap.add_argument("--batch", type=int, default=131072, help="per-GPU batch size (weak scaling)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-standin", action="store_true", help="skip the same-invocation stand-in arms")
    ap.add_argument("--no-compile", action="store_true", help="skip the torch.compile variant of the device stand-in")
    ap.add_argument("--no-actor", action="store_true", help="skip the init_spark -> fit_on_spark actor-mode run")
    ap.add_argument("--no-extras", action="store_true", help="skip the DLRM / BigMLP / collective records")
    ap.add_argument("--lean", action="store_true", help="only the two headline numbers (device + e2e)")
    ap.add_argument("--budget-s", type=float, default=float(os.environ.get("RAYDP_B200_BENCH_BUDGET_S", 300)),
                    help="wall-clock budget for the auxiliary phases (stand-ins, replica check, actor run, extras): a "
                         "phase that would not fit is skipped and listed under 'skipped'")
    return ap.parse_args()


from raydp_b200.utils.profiling import ClockSampler  # noqa: E402


def dist_setup(n):
    import torch
    import torch.distributed as dist
    if n == 1 and "RANK" not in os.environ:      # plain `python bench.py`: a one-rank SPMD world
        os.environ.update({"RANK": "0", "WORLD_SIZE": "1", "LOCAL_RANK": "0"})
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world != n:
        raise SystemExit(f"--gpus {n} but WORLD_SIZE={world}: launch with torch.distributed.run "
                         f"--nproc-per-node {n}")
    torch.cuda.set_device(local)
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    return rank, world, local


def max_over_ranks(value: float, world: int) -> float:
    import torch
    import torch.distributed as dist
    if world == 1:
        return value
    t = torch.tensor([value], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def barrier(world):
    import torch
    import torch.distributed as dist
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def synthetic_columns(rows: int, seed: int):
    """NYC-taxi-shaped rows: 20 float32 features (already engineered) + fare label."""
    import numpy as np
    rng = np.random.default_rng(seed)
    feats = rng.standard_normal((FEATURES, rows), dtype=np.float32)
    w = np.linspace(-1.0, 1.0, FEATURES, dtype=np.float32)
    label = (feats * w[:, None]).sum(0) * 2.0 + 10.0 + rng.standard_normal(rows, dtype=np.float32)
    return feats, label.astype(np.float32)


def _user_model():
    import torch
    from user_models import NYC_Model
    torch.manual_seed(1234)
    model = NYC_Model(FEATURES)
    assert not hasattr(model, "fused_spec")
    return model


def _device_pool(B, kpad, device, rank, pool_n=8, dtype=None):
    import torch
    dtype = dtype or torch.bfloat16
    g = torch.Generator(device=device).manual_seed(rank + 1)
    pool_x, pool_y = [], []
    for _ in range(pool_n):
        feats = torch.randn(B, FEATURES, device=device, generator=g)
        px = torch.zeros(B, kpad, dtype=dtype, device=device)
        px[:, :FEATURES] = feats.to(dtype)
        pool_x.append(px)
        pool_y.append(feats[:, :5].sum(1) * 2 + 10)
    return pool_x, pool_y


def _timed_loop(step, steps, warmup, world, local):
    """W warm-up steps, then K steps between barrier+synchronize, CUDA events, max over ranks."""
    import torch
    sampler = ClockSampler(local)
    sampler.start()
    sampler.wait_first_sample()
    for i in range(warmup):
        step(i)
    barrier(world)
    t_begin = time.time()
    t0 = time.perf_counter()
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    start.record()
    for i in range(steps):
        step(warmup + i)
    end.record()
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    t_end = time.time()
    barrier(world)
    clocks = sampler.stop(t_begin, t_end)
    return max_over_ranks(start.elapsed_time(end), world), max_over_ranks(wall * 1e3, world), clocks


# ------------------------------------------------------------------------------------------------
# ours: kernel-only number
# ------------------------------------------------------------------------------------------------
def _make_engine(args, rank, world, local, extra=None):
    import torch
    import torch.nn as nn
    from raydp_b200.torch.engine import make_engine
    from raydp_b200.torch.torch_metrics import TorchMetric
    from raydp_b200.train.session import TrainSession, _set_session
    device = torch.device("cuda", local)
    model = _user_model()
    optimizer = torch.optim.Adam(model.parameters(), lr=1e-3)
    config = {"batch_size": args.batch, "feature_columns": [f"f{i}" for i in range(FEATURES)],
              "label_column": "fare_amount", "feature_types": torch.float, "label_type": torch.float,
              "drop_last": True, "engine": "auto", "extra": dict(extra or {})}
    sess = TrainSession(rank, world, local, world, {}, device, "bench", "/tmp", "nccl")
    _set_session(sess)
    engine = make_engine(config, sess, model, optimizer, nn.SmoothL1Loss(), None,
                         TorchMetric(["MeanAbsoluteError", "MeanSquaredError"], None))
    return engine, device


def run_ours_device(args, rank, world, local):
    import torch
    from raydp_b200 import ops
    from raydp_b200.train.session import _set_session
    engine, device = _make_engine(args, rank, world, local)
    assert engine.name == "fused", f"the reference NYC_Model did not select the fused engine: {engine.name}"
    B = args.batch
    pool_x, pool_y = _device_pool(B, engine.fused.kpad, device, rank)
    x = torch.zeros(B, engine.fused.kpad, dtype=torch.bfloat16, device=device)
    y = torch.zeros(B, device=device)

    def step(i):
        x.copy_(pool_x[i % len(pool_x)])
        y.copy_(pool_y[i % len(pool_y)])
        engine._train_step(x, y)

    w = max(args.warmup, 3)
    for i in range(w):
        step(i)
    launches0 = ops.launch_count()
    ms, _, clocks = _timed_loop(step, args.steps, 0, world, local)
    launches = ops.launch_count() - launches0
    loss = float((engine.fused.accum[0] / engine.fused.accum[3]).item())
    engine.close()
    _set_session(None)
    return {"ms_per_step": ms / args.steps, "value": world * B * args.steps / (ms / 1e3),
            "clocks": clocks, "gpu_launches": int(launches), "final_loss": loss, "engine": engine.name}


def run_replica_check(args, rank, world, local, steps=6):
    """After `steps` identical steps: (a) every rank holds bit-identical parameters; (b) the parameters
    differ from a run whose gradients were reduced by NCCL (stock all_reduce) by at most rounding."""
    import torch
    import torch.distributed as dist
    from raydp_b200.train.session import _set_session
    finals = {}
    for how in ("p2p", "nccl"):
        engine, device = _make_engine(args, rank, world, local, extra={"allreduce": how, "cuda_graph": how == "p2p"})
        pool_x, pool_y = _device_pool(args.batch, engine.fused.kpad, device, rank, pool_n=2)
        for i in range(steps):
            engine._train_step(pool_x[i % 2], pool_y[i % 2])
        torch.cuda.synchronize()
        finals[how] = engine.fused.params.detach().clone()
        if getattr(engine, "group_reducers", None):
            kind = f"p2p per-layer x{len(engine.group_reducers)} (" + \
                   ("ll" if engine.group_reducers[0].ll else "one/two-shot") + "), overlapped with backward"
        else:
            kind = engine.reducer.kind if engine.reducer is not None else "none"
        finals[how + "_kind"] = kind
        engine.close()
        _set_session(None)
    mine = finals["p2p"]
    ref = mine.clone()
    dist.broadcast(ref, src=0)
    same = torch.tensor([1 if torch.equal(ref, mine) else 0], device="cuda")
    dist.all_reduce(same, op=dist.ReduceOp.MIN)
    diff = (finals["p2p"] - finals["nccl"]).abs().max()
    scale = finals["nccl"].abs().max()
    dist.all_reduce(diff, op=dist.ReduceOp.MAX)
    return {"ok": bool(same.item()) and float(diff.item()) < 1e-2 * max(1.0, float(scale.item())),
            "bit_identical_across_ranks": bool(same.item()), "steps": steps,
            "max_abs_diff_vs_nccl_reducer": float(diff.item()), "param_abs_max": float(scale.item()),
            "reducer": finals["p2p_kind"], "reference_reducer": finals["nccl_kind"]}


# ------------------------------------------------------------------------------------------------
# ours: end to end through TorchEstimator.fit (SPMD: every rank calls fit on its local shard)
# ------------------------------------------------------------------------------------------------
def run_ours_e2e(args, rank, world, local):
    import pyarrow as pa
    import torch
    import torch.nn as nn
    from raydp_b200 import data as rdata
    from raydp_b200 import runtime
    from raydp_b200.torch import TorchEstimator

    os.environ["RAYDP_B200_SPMD"] = "1"
    if not runtime.is_initialized():
        runtime.init(num_cpus=2, num_gpus=0)
    B = args.batch
    steps = max(args.steps, args.warmup)
    rows = steps * B
    feats, label = synthetic_columns(rows, seed=100 + rank)
    names = [f"f{i}" for i in range(FEATURES)]
    # a handful of blocks per rank, like partitions coming out of from_spark
    n_blocks = 4
    per = rows // n_blocks
    tables = []
    for b in range(n_blocks):
        lo, hi = b * per, (rows if b == n_blocks - 1 else (b + 1) * per)
        cols = {n: pa.array(feats[i, lo:hi]) for i, n in enumerate(names)}
        cols["fare_amount"] = pa.array(label[lo:hi])
        tables.append(pa.table(cols))
    ds = rdata.from_arrow(tables)
    del feats, label, tables
    model = _user_model()
    est = TorchEstimator(num_workers=world, resources_per_worker={"GPU": 1}, model=model,
                         optimizer=torch.optim.Adam(model.parameters(), lr=1e-3), loss=nn.SmoothL1Loss(),
                         feature_columns=names, feature_types=torch.float, label_column="fare_amount",
                         label_type=torch.float, batch_size=B, num_epochs=2, drop_last=True, shuffle=False,
                         metrics_name=["MeanAbsoluteError", "MeanSquaredError"], use_gpu=True,
                         spmd_local_shards=True)
    result = est.fit(ds)
    epochs = [h for h in result.metrics_history if "train_loss" in h]
    timed = epochs[-1]                       # epoch 0 = warm-up (>= W steps), epoch 1 = timed
    secs = max_over_ranks(max(timed["time_this_iter_s"], timed.get("device_time_s", 0.0)), world)
    k = timed["steps_this_iter"]
    return {"value": world * timed["samples_this_iter"] / secs, "unit": "samples/s",
            "h2d_bytes_per_step": int(timed["h2d_bytes_this_iter"] / k),
            "d2h_bytes_per_step": int(timed["d2h_bytes_this_iter"] / k),
            "steps": k, "ms_per_step": secs * 1e3 / k, "engine": timed["engine"],
            "train_loss": timed["train_loss"], "mode": "SPMD (one torchrun rank per GPU calls fit)"}


def run_actor_e2e(args, world):
    """The default user path, once, from rank 0 only: init_spark -> DataFrame -> fit_on_spark with
    TrainWorker actors.  Rows are generated by the executors (spark.range + withColumn(rand))."""
    import torch
    import torch.nn as nn
    import raydp_b200 as raydp
    from raydp_b200 import runtime
    from raydp_b200.sql.functions import col, rand
    from raydp_b200.torch import TorchEstimator
    os.environ["RAYDP_B200_SPMD"] = "0"
    os.environ.setdefault("RAYDP_B200_PG_TIMEOUT_S", "120")     # a worker group that cannot form fails in 2 minutes
    runtime.shutdown()
    n_exec = max(2, min(8, 2 * world))
    runtime.init(num_cpus=n_exec + world + 2, num_gpus=world)
    # BASELINE.json config 2 names a 100 M-row frame on 8 GPUs: 12.5 M rows (95 steps of 131072) per GPU, if the
    # shared-memory object store has room for it (84 bytes per row, x3 for the conversion's transient copies)
    steps_per_rank = 95
    import shutil
    free = shutil.disk_usage("/dev/shm").free
    if steps_per_rank * args.batch * world * (FEATURES + 1) * 4 * 3 > free:
        steps_per_rank = 16
    rows = steps_per_rank * args.batch * world
    t0 = time.perf_counter()
    spark = raydp.init_spark("bench-actor", n_exec, 1, "4GB")
    df = spark.range(0, rows, numPartitions=4 * n_exec)
    for i in range(FEATURES):
        df = df.withColumn(f"f{i}", (rand(i + 1) * 2 - 1).cast("float"))
    label = col("f0") * 3 + col("f1") * 2 - col("f2") + 5
    df = df.withColumn("fare_amount", label.cast("float")).drop("id")
    t_init = time.perf_counter() - t0
    model = _user_model()
    names = [f"f{i}" for i in range(FEATURES)]
    est = TorchEstimator(num_workers=world, resources_per_worker={"GPU": 1, "CPU": 1}, model=model,
                         optimizer=torch.optim.Adam(model.parameters(), lr=1e-3), loss=nn.SmoothL1Loss(),
                         feature_columns=names, feature_types=torch.float, label_column="fare_amount",
                         label_type=torch.float, batch_size=args.batch, num_epochs=3, drop_last=True, shuffle=False,
                         metrics_name=["MeanAbsoluteError", "MeanSquaredError"])
    from raydp_b200.spark.dataset import spark_dataframe_to_ray_dataset
    t0 = time.perf_counter()
    ds = spark_dataframe_to_ray_dataset(df)
    n = ds.count()
    t_conv = time.perf_counter() - t0
    t0 = time.perf_counter()
    result = est.fit(ds, max_retries=0)
    t_fit = time.perf_counter() - t0
    epochs = [h for h in result.metrics_history if "train_loss" in h]
    timed = epochs[-1]
    secs = max(timed["time_this_iter_s"], timed.get("device_time_s", 0.0))
    out = {"value": world * timed["samples_this_iter"] / secs, "unit": "samples/s (rank 0's last epoch x N)",
           "rows": n, "engine": timed["engine"], "steps_per_epoch": timed["steps_this_iter"],
           "ms_per_step": secs * 1e3 / max(1, timed["steps_this_iter"]),
           "init_spark_and_plan_s": t_init, "conversion_s": t_conv, "fit_s_incl_worker_start": t_fit,
           "executors": n_exec, "mode": "actors (init_spark -> spark.range/withColumn -> from_spark -> fit)",
           "train_loss": timed["train_loss"]}
    raydp.stop_spark()
    runtime.shutdown()
    os.environ["RAYDP_B200_SPMD"] = "1"
    return out


# ------------------------------------------------------------------------------------------------
# stand-ins: the reference loop on stock parts
# ------------------------------------------------------------------------------------------------
def _stock_model(local, world):
    import torch
    model = _user_model().to(torch.device("cuda", local))
    if world > 1:
        model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
    return model


def run_standin_device(args, rank, world, local, compiled=False):
    """Stock torch on device-resident batches: bf16 autocast, cuBLAS/cuDNN, DDP/NCCL, fused Adam, no host sync."""
    import torch
    import torch.nn as nn
    device = torch.device("cuda", local)
    model = _stock_model(local, world)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3, fused=True)
    crit = nn.SmoothL1Loss()
    B = args.batch
    pool_x, pool_y = _device_pool(B, FEATURES, device, rank, dtype=torch.float32)
    pool_y = [y.unsqueeze(1) for y in pool_y]
    fwd = model
    if compiled:
        fwd = torch.compile(model)

    def step(i):
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = fwd(pool_x[i % len(pool_x)])
        loss = crit(out.float(), pool_y[i % len(pool_y)])
        opt.zero_grad(set_to_none=True)
        loss.backward()
        opt.step()

    ms, _, clocks = _timed_loop(step, args.steps, max(args.warmup, 3), world, local)
    return {"ms_per_step": ms / args.steps, "value": world * B * args.steps / (ms / 1e3), "clocks": clocks,
            "what": "stock torch, device-resident batches: bf16 autocast + cuBLAS + DDP/NCCL + fused Adam"
                    + (" + torch.compile" if compiled else " (eager)")}


def run_standin(args, rank, world, local):
    """The reference's loop (torch/estimator.py:272-293) minimally patched for GPU: to_torch-style CPU
    batch assembly, blocking H2D, bf16 autocast, per-step loss.item(), DDP/NCCL, torch.optim.Adam."""
    import numpy as np
    import torch
    import torch.nn as nn
    device = torch.device("cuda", local)
    model = _stock_model(local, world)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    crit = nn.SmoothL1Loss()
    B = args.batch
    w = max(args.warmup, 3)
    steps = args.steps + w
    feats, label = synthetic_columns(steps * B, seed=100 + rank)
    cols = [feats[i] for i in range(FEATURES)]

    def batch(i):      # to_torch semantics: per-column numpy -> [B, F] float32 tensor on the CPU
        lo, hi = i * B, (i + 1) * B
        arr = np.empty((B, FEATURES), dtype=np.float32)
        for j, c in enumerate(cols):
            arr[:, j] = c[lo:hi]
        return torch.from_numpy(arr), torch.from_numpy(label[lo:hi].reshape(-1, 1).copy())

    def step(i):
        inputs, targets = batch(i)
        inputs, targets = inputs.to(device), targets.to(device)
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = model(inputs)
        loss = crit(out.float(), targets)
        total = loss.item()
        opt.zero_grad()
        loss.backward()
        opt.step()
        return total

    ms, wall_ms, clocks = _timed_loop(step, args.steps, w, world, local)
    secs = max(wall_ms, ms) / 1e3
    value = world * B * args.steps / secs
    return {"ms_per_step": secs * 1e3 / args.steps, "value": value, "clocks": clocks, "gpu_launches": 0,
            "what": "reference loop on stock parts: CPU batch assembly + blocking H2D + loss.item() per step",
            "e2e": {"value": value, "unit": "samples/s", "h2d_bytes_per_step": B * (FEATURES + 1) * 4,
                    "d2h_bytes_per_step": 4}}


_T0 = 0.0


def _guard(errors, name, fn, *a, **kw):
    """Auxiliary phases must never cost the headline line: record the failure and go on."""
    try:
        return fn(*a, **kw)
    except BaseException as exc:  # noqa: BLE001
        if isinstance(exc, (KeyboardInterrupt, SystemExit)):
            raise
        errors[name] = f"{type(exc).__name__}: {exc}"
        traceback.print_exc(file=sys.stderr)
        return None


def main():
    global _T0
    _T0 = time.perf_counter()
    args = parse_args()
    if args.impl == "reference":
        print(json.dumps({"impl": "reference", "unavailable":
                          "pang-wu/raydp needs Maven+JDK to build its jars and ray+pyspark at import; "
                          "none are in this offline image (pip install of /root/reference fails, see DESIGN.md)"}))
        return 0
    if args.lean:
        args.no_standin = args.no_actor = args.no_extras = True
    rank, world, local = dist_setup(args.gpus)
    import torch
    config = {"model": "NYC-taxi MLP 20-256-128-64-16-1 (Linear-ReLU-BatchNorm1d), SmoothL1, Adam; the reference "
                       "example's own NYC_Model class, unchanged",
              "global_batch": args.batch * world, "per_gpu_batch": args.batch, "seq_len": None,
              "parallelism": f"dp{world}", "features": FEATURES,
              "l2_policy": "per-step working set (activations+gradients > 150 MB) exceeds the 126 MB L2; "
                           "inputs rotate over 8 device buffers (kernel number) / stream from host (e2e)"}
    if args.impl == "standin":
        out = run_standin(args, rank, world, local)
        line = {"impl": "standin", "metric": METRIC, "value": out["value"], "unit": "samples/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": out["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16 autocast", "data": "synthetic", "config": config,
                "clocks": out["clocks"], "e2e": out["e2e"], "gpu_launches": 0}
    else:
        errors = {}
        dev = run_ours_device(args, rank, world, local)
        e2e = None if args.no_e2e else run_ours_e2e(args, rank, world, local)
        line = {"impl": "ours", "metric": METRIC, "value": dev["value"], "unit": "samples/s",
                "n_gpus": world, "steps": args.steps, "warmup": max(args.warmup, 3),
                "ms_per_step": dev["ms_per_step"], "higher_is_better": True, "scaling": "weak",
                "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "config": config,
                "clocks": dev["clocks"], "e2e": e2e, "gpu_launches": dev["gpu_launches"],
                "final_loss": dev["final_loss"], "engine": dev["engine"]}
        torch.cuda.empty_cache()
        # ---- auxiliary phases, most important first, inside a wall-clock budget (same decision on every rank)
        t_start = _T0
        phase_s, skipped = {}, []

        def fits(name, estimate):
            elapsed = time.perf_counter() - t_start
            if world > 1:
                import torch.distributed as dist
                t = torch.tensor([elapsed], device=torch.device("cuda", local))
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                elapsed = float(t.item())
            if elapsed + estimate > args.budget_s:
                skipped.append(name)
                return False
            return True

        def phase(name, fn, *a, **kw):
            t0 = time.perf_counter()
            out = _guard(errors, name, fn, *a, **kw)
            phase_s[name] = round(time.perf_counter() - t0, 1)
            torch.cuda.empty_cache()
            return out

        standin = {}
        if not args.no_standin:
            if fits("standin_device_eager", 20):
                standin["device_eager"] = phase("standin_device_eager", run_standin_device, args, rank, world, local)
            if fits("standin_e2e", 40):
                se = phase("standin_e2e", run_standin, args, rank, world, local)
                standin["e2e"] = {k: se[k] for k in ("ms_per_step", "value", "clocks", "what")} if se else None
        if world > 1 and fits("replica_check", 15):
            line["replica_check"] = phase("replica_check", run_replica_check, args, rank, world, local)
        if not args.no_actor and fits("e2e_actor", 90):
            # rank 0 drives an actor-mode fit on all GPUs; the other ranks wait on the host (no kernel
            # of theirs may spin on the GPUs the actors are using)
            store = None
            if world > 1:
                import torch.distributed as dist
                torch.cuda.synchronize()
                dist.barrier()
                torch.cuda.synchronize()
                store = dist.distributed_c10d._get_default_store()
            if rank == 0:
                line["e2e_actor"] = phase("e2e_actor", run_actor_e2e, args, world)
                if store is not None:
                    store.set("bench_actor_phase_done", "1")
            elif store is not None:
                import datetime
                store.wait(["bench_actor_phase_done"], datetime.timedelta(minutes=12))
        if not args.no_extras:
            sys.path.insert(0, os.path.join(ROOT, "bench"))
            import extras
            extra = {}
            if fits("extra_dlrm", 40):
                extra["dlrm"] = phase("extra_dlrm", extras.dlrm_record, rank, world, local)
            if world > 1 and fits("extra_allreduce", 40):
                extra["allreduce"] = phase("extra_allreduce", extras.allreduce_record, rank, world, local)
            if world > 1 and fits("extra_alltoall", 20):
                extra["alltoall"] = phase("extra_alltoall", extras.alltoall_record, rank, world, local)
            if fits("extra_bigmlp", 60):
                extra["bigmlp"] = phase("extra_bigmlp", extras.bigmlp_record, rank, world, local)
            line["extra"] = extra
        if not args.no_standin and not args.no_compile and fits("standin_device_compiled", 60):
            standin["device_compiled"] = phase("standin_device_compiled", run_standin_device, args, rank, world, local,
                                               compiled=True)
        if standin:
            line["standin"] = standin
            dev_best = max([s["value"] for s in (standin.get("device_eager"), standin.get("device_compiled")) if s],
                           default=None)
            detail = {"device": dev["value"] / dev_best if dev_best else None,
                      "e2e": (e2e["value"] / standin["e2e"]["value"]) if (e2e and standin.get("e2e")) else None,
                      "note": "ours / reference-equivalent stand-in on stock torch+NCCL+cuBLAS, same invocation; "
                              "the reference itself publishes no number and cannot be installed here"}
            line["vs_baseline"] = detail["device"]
            line["vs_baseline_detail"] = detail
        line["phase_seconds"] = phase_s
        if skipped:
            line["skipped"] = {"phases": skipped, "why": f"auxiliary-phase budget of {args.budget_s:.0f} s (--budget-s)"}
        if errors:
            line["errors"] = errors
    if rank == 0:
        print(json.dumps(line), flush=True)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    try:
        from raydp_b200 import runtime
        runtime.shutdown()
    except Exception:  # noqa: BLE001
        pass
    return 0


if __name__ == "__main__":
    sys.exit(main())
