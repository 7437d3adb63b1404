"""Where does the forward GEMM of the flagship MLP spend its time?  Times rdp_gemm_tn on the step's own shapes
with the epilogue features switched on one by one (CUDA events, 20 iterations after 5 warm-ups, operands far
larger than L2 and rotated).  Diagnostic, not a benchmark record."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

from raydp_b200 import ops  # noqa: E402


def timed(fn, iters=20, warm=5):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(iters):
        fn()
    e.record()
    torch.cuda.synchronize()
    return s.elapsed_time(e) / iters * 1e3


def main():
    dev = torch.device("cuda", 0)
    B = 131072
    for K, N in ((24, 256), (32, 256), (64, 256), (256, 128), (128, 64), (64, 16)):
        xs = [torch.randn(B, K, device=dev).to(torch.bfloat16) for _ in range(3)]
        w = torch.randn(N, K, device=dev).to(torch.bfloat16)
        bias = torch.randn(N, device=dev)
        outs = [torch.empty(B, N, dtype=torch.bfloat16, device=dev) for _ in range(3)]
        stats = torch.zeros(2, N, device=dev)
        aux = torch.randn(B, N, device=dev).to(torch.bfloat16)
        i = [0]

        def run(**kw):
            i[0] += 1
            ops.gemm_tn(xs[i[0] % 3], w, outs[i[0] % 3], **kw)
        mb = (B * K * 2 + B * N * 2) / 1e6
        row = {"plain": timed(lambda: run()), "bias+relu": timed(lambda: run(bias=bias, relu=True)),
               "bias+relu+stats1": timed(lambda: run(bias=bias, relu=True, stats_mode=1, stats=stats))}
        if N % 64 == 0:
            row["aux mask"] = timed(lambda: run(aux=aux, aux_relu_mask=True))
            row["aux stats2"] = timed(lambda: run(aux=aux, stats_mode=2, stats=stats))
        copy = timed(lambda: outs[0].copy_(outs[1]))
        print(f"K={K:4d} N={N:4d} traffic {mb:6.1f} MB  " + "  ".join(f"{k} {v:6.1f}us ({mb / v / 1e3 * 1e3:5.2f} TB/s)"
              for k, v in row.items()) + f"  | copy of the output {copy:5.1f}us", flush=True)


if __name__ == "__main__":
    main()
