"""One training step of an ARBITRARY user module on the generic GPU engine's parts, for a kernel launch list.

    python bench/generic_step.py
    ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
        --log-file gpurun_out/generic_step_launches.csv python bench/generic_step.py

The module (Linear / GELU / Dropout / LayerNorm / a residual add -- nothing the MLP or DLRM lowerings accept) goes
through exactly what `GenericGpuEngine` does with it: `swap_linears` (every nn.Linear -> FusedLinear on the tcgen05
GEMMs), bf16 autocast for the rest, `FusedDataParallel.backward_and_step` (fused Adam + bf16 operand shadow).  The
launch list of the profiled step is the evidence that no cuBLAS (`nvjet`, `cutlass`) or NCCL kernel runs on this
path: summary in profiles/generic_step_launches.md.
"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))

import torch  # noqa: E402
import torch.nn as nn  # noqa: E402


class Block(nn.Module):
    def __init__(self, width, hidden):
        super().__init__()
        self.norm = nn.LayerNorm(width)
        self.up = nn.Linear(width, hidden)
        self.act = nn.GELU()
        self.drop = nn.Dropout(0.1)
        self.down = nn.Linear(hidden, width)

    def forward(self, x):
        return x + self.down(self.drop(self.act(self.up(self.norm(x)))))


class UserModel(nn.Module):
    def __init__(self, features=40, width=512, hidden=2048, depth=3):
        super().__init__()
        self.inp = nn.Linear(features, width)
        self.blocks = nn.Sequential(*[Block(width, hidden) for _ in range(depth)])
        self.out = nn.Linear(width, 1)

    def forward(self, x):
        return self.out(self.blocks(self.inp(x)))


def main():
    from raydp_b200.ops.linear import swap_linears
    from raydp_b200.parallel import FusedDataParallel
    dev = torch.device("cuda", 0)
    torch.manual_seed(0)
    model = UserModel().to(dev)
    n = swap_linears(model)
    opt = torch.optim.Adam(model.parameters(), lr=1e-3)
    dp = FusedDataParallel(model, opt, dev)
    crit = nn.SmoothL1Loss()
    B = 16384
    x = torch.randn(B, 40, device=dev)
    y = torch.randn(B, 1, device=dev)

    def step():
        with torch.autocast("cuda", dtype=torch.bfloat16):
            out = dp(x)
        loss = crit(out.float(), y)
        dp.backward_and_step(loss)
        return loss
    for _ in range(3):
        step()
    torch.cuda.synchronize()
    torch.cuda.profiler.start()
    loss = step()
    torch.cuda.synchronize()
    torch.cuda.profiler.stop()
    print(f"fused linears: {n}; loss {float(loss):.4f}", flush=True)


if __name__ == "__main__":
    main()
