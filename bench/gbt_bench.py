"""Histogram gradient boosting: one boosting round on the CPU (numpy bincount) vs on the GPU (`csrc/kernels/gbt.cu`).

    python bench/gbt_bench.py [--rows 1000000] [--features 28] [--depth 6] [--rounds 5]

Synthetic regression table (HIGGS-shaped: 28 features); "step" = one boosting iteration (one tree).  The GPU number is
wall time per round around `hist_gbt.train` with `tree_method="gpu_hist"` (it includes the host-side split search and
the per-round upload of the gradients); the kernels' own share is reported from CUDA events."""
import argparse
import json
import os
import sys
import time

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import numpy as np  # noqa: E402
import torch  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=1_000_000)
    ap.add_argument("--features", type=int, default=28)
    ap.add_argument("--depth", type=int, default=6)
    ap.add_argument("--rounds", type=int, default=5)
    ap.add_argument("--cpu-rounds", type=int, default=2)
    args = ap.parse_args()
    from raydp_b200 import ops
    from raydp_b200.xgboost import hist_gbt
    rng = np.random.default_rng(0)
    x = rng.standard_normal((args.rows, args.features)).astype(np.float32)
    y = (x[:, 0] * 2 - x[:, 1] ** 2 + 0.5 * x[:, 2] * x[:, 3] + 0.1 * rng.standard_normal(args.rows)).astype(np.float64)
    params = {"objective": "reg:squarederror", "max_depth": args.depth, "eta": 0.3}
    out = {"rows": args.rows, "features": args.features, "max_depth": args.depth, "max_bin": 256}
    stamps = []
    hist_gbt.train(dict(params, tree_method="gpu_hist"), x[:50000], y[:50000], 1)          # warm-up (context, kernels)
    t0 = time.perf_counter()
    _, hist = hist_gbt.train(dict(params, tree_method="gpu_hist"), x, y, args.rounds,
                             callback=lambda r, m, met: stamps.append(time.perf_counter()))
    per_round = np.diff([stamps[0]] + stamps[1:]) if len(stamps) > 1 else [stamps[0] - t0]
    out["gpu_hist"] = {"s_per_round": float(np.median(per_round)), "setup_s_incl_binning": stamps[0] - t0 - float(np.median(per_round)),
                       "rmse": hist["train"]["rmse"][-1]}
    # kernels alone: one level-0 histogram and one deepest-level histogram + a partition pass
    dev = torch.device("cuda", 0)
    cuts = hist_gbt._global_cuts(x, 256, hist_gbt._Comm())
    binned, n_bins = hist_gbt._bin(x, cuts)
    bt = torch.from_numpy(np.ascontiguousarray(binned.T).view(np.int16)).to(dev)
    g = torch.randn(args.rows, device=dev)
    h = torch.rand(args.rows, device=dev)
    for slots in (1, 1 << (args.depth - 1)):
        slot = torch.randint(0, slots, (args.rows,), device=dev, dtype=torch.int32)
        hd = torch.zeros(slots, 2, args.features, n_bins, dtype=torch.float64, device=dev)
        for _ in range(3):
            ops.gbt_hist(bt, g, h, slot, hd)
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(10):
            ops.gbt_hist(bt, g, h, slot, hd)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / 10
        out[f"hist_kernel_{slots}_nodes"] = {"ms": ms, "G_updates_per_s": 2 * args.rows * args.features / ms / 1e6,
                                              "GBps_read": args.rows * (args.features * 2 + 12) / ms / 1e6}
    stamps = []
    t0 = time.perf_counter()
    _, hist = hist_gbt.train(dict(params, device="cpu"), x, y, args.cpu_rounds,
                             callback=lambda r, m, met: stamps.append(time.perf_counter()))
    per_round = np.diff(stamps) if len(stamps) > 1 else [stamps[0] - t0]
    out["cpu_hist"] = {"s_per_round": float(np.median(per_round)), "rmse": hist["train"]["rmse"][-1]}
    out["gpu_over_cpu"] = out["cpu_hist"]["s_per_round"] / out["gpu_hist"]["s_per_round"]
    print(json.dumps(out), flush=True)


if __name__ == "__main__":
    main()
