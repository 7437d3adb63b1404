"""GEMM micro-benchmark at the large-MLP shapes: CTA-pair tcgen05 kernels vs cuBLAS (torch.matmul).

    python bench/gemm_bench.py [--m 8192 --n 16384 --k 8192] [--iters 10]

Prints one JSON line per op (fwd/dgrad-shaped TN GEMM, wgrad) with TFLOP/s for the forced 1-CTA
path, the CTA-pair path and cuBLAS.  Inputs (~400 MB per GEMM) exceed the 126 MB L2.
"""
import argparse
import json
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402


def timed(fn, iters):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--m", type=int, default=8192)
    ap.add_argument("--n", type=int, default=16384)
    ap.add_argument("--k", type=int, default=8192)
    ap.add_argument("--iters", type=int, default=10)
    args = ap.parse_args()
    from raydp_b200 import ops
    M, N, K = args.m, args.n, args.k
    dev = torch.device("cuda")
    a = (torch.randn(M, K, device=dev) * 0.1).to(torch.bfloat16)
    b = (torch.randn(N, K, device=dev) * 0.1).to(torch.bfloat16)
    out = torch.empty(M, N, dtype=torch.bfloat16, device=dev)
    flops = 2.0 * M * N * K

    def compare(variants, rounds=3):
        """Interleave the variants so clock/thermal drift hits all of them alike; report the median."""
        got = {k: [] for k in variants}
        for _ in range(rounds):
            for name, fn in variants.items():
                got[name].append(flops / timed(fn, args.iters) / 1e9)
        return {k + "_tflops": round(sorted(v)[len(v) // 2], 1) for k, v in got.items()}

    res = {"op": "gemm_tn", "M": M, "N": N, "K": K}
    def dyn512():
        ops.set_dynamic_tiles(True)
        ops.gemm2_tn(a, b, out, mh=2)
        ops.set_dynamic_tiles(False)

    res.update(compare({"pair256": lambda: ops.gemm2_tn(a, b, out, mh=1),
                        "pair512": lambda: ops.gemm2_tn(a, b, out, mh=2),
                        "pair512_clc": dyn512,
                        "cublas": lambda: torch.matmul(a, b.t(), out=out)}))
    ref = torch.matmul(a, b.t())
    ops.gemm2_tn(a, b, out)
    res["max_abs_diff_vs_cublas"] = float((out.float() - ref.float()).abs().max())
    print(json.dumps(res), flush=True)
    del ref
    # wgrad: dW[N, K] += g[M, N]^T x[M, K]
    g = (torch.randn(M, N, device=dev) * 0.1).to(torch.bfloat16)
    w = torch.zeros(N, K, device=dev)
    wb = torch.empty(N, K, dtype=torch.bfloat16, device=dev)
    res = {"op": "gemm_wgrad", "rows": M, "Mw": N, "Nw": K}
    res.update(compare({"pair256": lambda: ops.gemm2_wgrad(g, a, w, mh=1),
                        "pair512": lambda: ops.gemm2_wgrad(g, a, w, mh=2),
                        "cublas_bf16_out": lambda: torch.matmul(g.t(), a, out=wb)}))
    w.zero_()
    ops.gemm2_wgrad(g, a, w)
    res["max_rel_diff_vs_cublas"] = float((w - torch.matmul(g.t(), a).float()).abs().max() / w.abs().max())
    print(json.dumps(res), flush=True)


if __name__ == "__main__":
    main()
