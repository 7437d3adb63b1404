"""Eager (no CUDA graph) fused training steps for kernel-level profiling with ncu.

    python bench/profile_step.py --batch 131072 --steps 5
    ncu --metrics gpu__time_duration.sum --clock-control none -s 120 -c 60 --csv \
        --log-file gpurun_out/launches.csv python bench/profile_step.py
"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402

from raydp_b200 import ops  # noqa: E402
from raydp_b200.models import NYCTaxiMLP  # noqa: E402
from raydp_b200.ops.fused_mlp import FusedMLP  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--batch", type=int, default=131072)
    ap.add_argument("--steps", type=int, default=5)
    args = ap.parse_args()
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = NYCTaxiMLP(20)
    fused = FusedMLP(model.fused_spec(), args.batch, dev)
    fused.load_state_dict({k: v.to(dev) for k, v in model.state_dict().items()})
    x = fused.pack_input(torch.randn(args.batch, 20, device=dev))
    y = torch.randn(args.batch, device=dev)
    m, v = torch.zeros_like(fused.params), torch.zeros_like(fused.params)
    start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for step in range(1, args.steps + 1):
        if step == args.steps:
            torch.cuda.synchronize()
            start.record()
        fused.train_step(x, y)
        ops.adam_step(fused.params, fused.grads, m, v, 1e-3, 0.9, 0.999, 1e-8, 0.0, step)
    end.record()
    torch.cuda.synchronize()
    print(f"last step {start.elapsed_time(end):.3f} ms, launches/step "
          f"{ops.launch_count() // args.steps}")


if __name__ == "__main__":
    main()
