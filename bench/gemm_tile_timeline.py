"""Per-tile timeline of the persistent tcgen05 GEMM (CTA 0), and event timings for operand row pitches.

    RDP_GEMM_DBG=3 python bench/gemm_tile_timeline.py 2> gpurun_out/gemm_tile_timeline.txt

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
    dev = torch.device("cuda", 0)
    B, N = 131072, 256
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    rows = []
    for K in (24, 32, 64):
        x = torch.randn(B, K, device=dev).to(torch.bfloat16)
        w = torch.randn(N, K, device=dev).to(torch.bfloat16)
        out = torch.empty(B, N, dtype=torch.bfloat16, device=dev)
        for _ in range(3):
            ops.gemm_tn(x, w, out)
        torch.cuda.synchronize()
        ts = []
        for _ in range(20):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record()
            ops.gemm_tn(x, w, out)
            b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b) * 1e3)
        ts.sort()
        ref = x.float() @ w.float().t()
        err = ((out.float() - ref).norm() / ref.norm()).item()
        rows.append({"K": K, "row_pitch_bytes": 2 * K, "us_median": round(ts[len(ts) // 2], 2),
                     "us_min": round(ts[0], 2), "rel_err_vs_fp32": err,
                     "GBps_out": round(B * N * 2 / (ts[len(ts) // 2] * 1e-6) / 1e9, 1)})
        print(json.dumps(rows[-1]), flush=True)


if __name__ == "__main__":
    main()
