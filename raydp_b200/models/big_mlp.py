"""Large MLP of the trainer benchmark (``BASELINE.json`` config 4: "1B-param MLP, fused all-reduce
overlapped with backward"; SURVEY.md section 2.6 workload C)."""
from __future__ import annotations

import torch
import torch.nn as nn


class BigMLP(nn.Module):
    """``depth`` blocks of Linear(width, hidden) + ReLU + Linear(hidden, width); default ~1.07 B params."""

    def __init__(self, width: int = 8192, hidden: int = 16384, depth: int = 4, fused: bool = True):
        super().__init__()
        from raydp_b200.ops.linear import FusedLinear
        layers = []
        for _ in range(depth):
            if fused:
                layers += [FusedLinear(width, hidden, relu=True), FusedLinear(hidden, width)]
            else:
                layers += [nn.Linear(width, hidden), nn.ReLU(), nn.Linear(hidden, width)]
        self.layers = nn.Sequential(*layers)
        self.head = nn.Linear(width, 1)

    def forward(self, x):
        return self.head(self.layers(x).float())
