"""Per-tile timeline of the persistent tcgen05 GEMM (CTA 0), and event timings for operand row pitches.

    python bench/gemm_tile_timeline.py > gpurun_out/gemm_pitch.jsonl
    RDP_GEMM_DBG=1 python bench/gemm_tile_timeline.py --k 64 --mode 1 2> gpurun_out/gemm_tile_timeline.txt

With RDP_GEMM_DBG=<n> the library stamps %globaltimer at fixed points of the first n launches (producer reaches
tile / loads issued / MMA warp reaches tile / accumulator free / MMAs issued / epilogue stages ...) and prints
them to stderr.  The second part times the first-layer GEMM of the NYC model (B=131072, N=256) with the input
rows at 48, 64 and 128 bytes pitch -- the measurement behind ``kpad = round_up(in_features, 32)`` in
raydp_b200/ops/fused_mlp.py.  Summary: profiles/gemm_tile_timeline.md.
"""
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from raydp_b200 import ops  # noqa: E402


def main():
    import argparse
    ap = argparse.ArgumentParser()
    ap.add_argument("--k", type=int, default=0, help="only this K (default: 24, 32, 64)")
    ap.add_argument("--mode", type=int, default=-1, help="0 plain, 1 with column statistics (default: both)")
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    B, N = 131072, 256
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    for K in ((args.k,) if args.k else (24, 32, 64)):
        x = torch.randn(B, K, device=dev).to(torch.bfloat16)
        w = torch.randn(N, K, device=dev).to(torch.bfloat16)
        out = torch.empty(B, N, dtype=torch.bfloat16, device=dev)
        stats = torch.zeros(2, N, device=dev)
        bias = torch.zeros(N, device=dev)
        for mode in ((args.mode,) if args.mode >= 0 else (0, 1)):     # plain store / + per-column statistics by extra MMAs
            kw = dict(bias=bias, relu=True, stats_mode=1, stats=stats) if mode else {}
            for _ in range(3):
                ops.gemm_tn(x, w, out, **kw)
            torch.cuda.synchronize()
            ts = []
            for _ in range(20):
                flush.zero_()
                a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                a.record()
                ops.gemm_tn(x, w, out, **kw)
                b.record()
                torch.cuda.synchronize()
                ts.append(a.elapsed_time(b) * 1e3)
            ts.sort()
            ref = x.float() @ w.float().t()
            if mode:
                ref = ref.clamp_min(0)
            err = ((out.float() - ref).norm() / ref.norm()).item()
            rows.append({"K": K, "row_pitch_bytes": 2 * K, "stats": bool(mode), "us_median": round(ts[len(ts) // 2], 2),
                         "us_min": round(ts[0], 2), "rel_err_vs_fp32": err,
                         "GBps_out": round(B * N * 2 / (ts[len(ts) // 2] * 1e-6) / 1e9, 1)})
            print(json.dumps(rows[-1]), flush=True)


if __name__ == "__main__":
    main()
