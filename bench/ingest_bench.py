"""K5 measurements: where does the end-to-end step lose time against the device-resident step?

    python bench/ingest_bench.py --steps 40                      # 1 GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 \
        --master-port 29547 bench/ingest_bench.py --steps 40

Per run (all device-timed, max over ranks, each with its own clock sample):

* ``h2d_raw``        one large ``cudaMemcpyAsync`` per column buffer straight from the pinned shared-memory
                     blocks: the PCIe ceiling of this box at this N;
* ``loader_only``    ``ArrowGpuLoader`` alone (staging copies + pack kernel, nothing consuming the batches);
* ``step_device``    the fused training step on device-resident batches (no loader);
* ``step_e2e``       loader + training step, as ``TorchEstimator.fit`` runs them;
* ``host_us_per_batch``  CPU time the loader + step spend ENQUEUEING one batch (no synchronisation inside).

``e2e / device`` near 1 means the copies hide under the step; ``loader_only`` above ``step_device`` means the
host->device path itself is the bound at this N.
"""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

import bench as hb  # noqa: E402  (bench.py at the repository root)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--steps", type=int, default=40)
    ap.add_argument("--batch", type=int, default=131072)
    ap.add_argument("--gpus", type=int, default=int(os.environ.get("WORLD_SIZE", "1")))
    ap.add_argument("--json", default=None)
    ap.add_argument("--timeline", default=None, help="write a chrome trace of 8 end-to-end steps (torch profiler; "
                                                     "the steps under the profiler are not timed)")
    args = ap.parse_args()
    args.warmup = 5
    rank, world, local = hb.dist_setup(args.gpus)
    import pyarrow as pa
    from raydp_b200 import data as rdata
    from raydp_b200 import ops, runtime, store
    from raydp_b200.ops.ingest import ArrowGpuLoader
    from raydp_b200.utils.affinity import bind_to_gpu
    from raydp_b200.utils.profiling import ClockSampler
    cpus = bind_to_gpu(local, local, world)
    os.environ["RAYDP_B200_SPMD"] = "1"
    runtime.init(num_cpus=2, num_gpus=0)
    dev = torch.device("cuda", local)
    B, steps = args.batch, args.steps
    rows = (steps + args.warmup + (8 if args.timeline else 0)) * B
    feats, label = hb.synthetic_columns(rows, seed=100 + rank)
    names = [f"f{i}" for i in range(hb.FEATURES)]
    per = rows // 4
    tables = []
    for b in range(4):
        lo, hi = b * per, (rows if b == 3 else (b + 1) * per)
        cols = {n: pa.array(feats[i, lo:hi]) for i, n in enumerate(names)}
        cols["fare_amount"] = pa.array(label[lo:hi])
        tables.append(pa.table(cols))
    ds = rdata.from_arrow(tables)
    del feats, label, tables
    out = {"n_gpus": world, "per_gpu_batch": B, "steps": steps, "bytes_per_batch": B * 21 * 4,
           "cpu_affinity": cpus if cpus is None else [cpus[0], cpus[-1]]}

    def timed(fn, n, warm):
        sampler = ClockSampler(local)
        sampler.start()
        sampler.wait_first_sample()
        for _ in range(warm):
            fn()
        hb.barrier(world)
        t_begin = time.time()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        s.record()
        for _ in range(n):
            fn()
        host = time.perf_counter() - t0
        e.record()
        torch.cuda.synchronize()
        t_end = time.time()
        hb.barrier(world)
        return hb.max_over_ranks(s.elapsed_time(e), world) / n, hb.max_over_ranks(host * 1e3, world) / n, \
            sampler.stop(t_begin, t_end)

    # ---- raw H2D from the pinned blocks -------------------------------------------------------
    loader = ArrowGpuLoader(ds, names, "fare_amount", B, True, dev, ld=24)
    loader._prepare()
    out["pinned_objects"], out["pin_failures"] = store.pinned_objects, store.pin_failures
    ch = loader._chunks[0]
    stage = loader._stage[0]
    side = torch.cuda.Stream()

    def raw():
        with torch.cuda.stream(side):
            ops.ingest_stage([stage.data_ptr() + o for o in ch.dst_off], ch.src, ch.nbytes, side)
        torch.cuda.current_stream().wait_stream(side)
    ms, _, clocks = timed(raw, 20, 3)
    out["h2d_raw"] = {"ms_per_chunk": ms, "chunk_bytes": ch.bytes_total, "GBs_per_gpu": ch.bytes_total / ms / 1e6,
                      "GBs_total": world * ch.bytes_total / ms / 1e6, "copies_per_chunk": ch.n_copies, "clocks": clocks}

    # ---- loader alone ----------------------------------------------------------------------------
    it = iter(loader)

    def load_only():
        next(it)
    ms, host, clocks = timed(load_only, steps, args.warmup)
    out["loader_only"] = {"ms_per_batch": ms, "GBs_per_gpu": out["bytes_per_batch"] / ms / 1e6,
                          "host_us_per_batch": host * 1e3, "clocks": clocks}
    del it

    # ---- training step: device-resident vs end to end --------------------------------------------
    targs = argparse.Namespace(batch=B, steps=steps, warmup=args.warmup)
    engine, _ = hb._make_engine(targs, rank, world, local)
    pool_x, pool_y = hb._device_pool(B, engine.fused.kpad, dev, rank, pool_n=4)
    x = torch.zeros_like(pool_x[0])
    y = torch.zeros_like(pool_y[0])
    state = {"i": 0}

    def step_device():
        i = state["i"] = state["i"] + 1
        x.copy_(pool_x[i % 4])
        y.copy_(pool_y[i % 4])
        engine._train_step(x, y)
    ms, host, clocks = timed(step_device, steps, args.warmup)
    out["step_device"] = {"ms_per_step": ms, "host_us_per_step": host * 1e3, "clocks": clocks}
    loader2 = ArrowGpuLoader(ds, names, "fare_amount", B, True, dev, ld=engine.fused.kpad)
    it2 = iter(loader2)

    def step_e2e():
        xb, yb = next(it2)
        engine._train_step(xb, yb)            # publishes the 32-byte result to pinned host memory itself
    ms, host, clocks = timed(step_e2e, steps, args.warmup)
    out["step_e2e"] = {"ms_per_step": ms, "host_us_per_batch": host * 1e3, "clocks": clocks}
    out["e2e_over_device"] = out["step_device"]["ms_per_step"] / out["step_e2e"]["ms_per_step"]
    if args.timeline:                       # every rank steps (collectives); rank 0 records
        import contextlib
        from torch.profiler import ProfilerActivity, profile
        hb.barrier(world)
        with (profile(activities=[ProfilerActivity.CUDA]) if rank == 0 else contextlib.nullcontext()) as prof:
            for _ in range(8):
                step_e2e()
            torch.cuda.synchronize()
        if rank == 0:
            prof.export_chrome_trace(args.timeline)
    engine.close()
    if rank == 0:
        print(json.dumps(out), flush=True)
        if args.json:
            with open(args.json, "w") as fh:
                json.dump(out, fh, indent=1)
    if world > 1:
        import torch.distributed as dist
        dist.barrier()
        dist.destroy_process_group()
    runtime.shutdown()


if __name__ == "__main__":
    main()
