"""Tabular MLPs with the reference example's layer order (Linear -> ReLU -> BatchNorm1d).

Parity: ``NYC_Model`` of reference ``examples/pytorch_nyctaxi.py:69-92`` (20 -> 256 -> 128 -> 64 ->
16 -> 1).  Parameter names (``fc1`` ... ``fc5``, ``bn1`` ... ``bn4``) are the same, so state_dicts
are interchangeable with the user's own copy of that model.  The module itself is plain torch
(it is what runs on CPU workers and what the numerics tests compare against);
``fused_spec()`` tells the GPU engine that the whole step can run on the fused kernels.
"""
from __future__ import annotations

from typing import Sequence

import torch
import torch.nn as nn
import torch.nn.functional as F


class TabularMLP(nn.Module):
    def __init__(self, in_features: int, hidden: Sequence[int] = (256, 128, 64, 16), out_features: int = 1,
                 batch_norm: bool = True):
        super().__init__()
        self.in_features = in_features
        self.hidden = tuple(hidden)
        self.out_features = out_features
        self.batch_norm = batch_norm
        dims = [in_features] + list(hidden)
        for i in range(len(hidden)):
            setattr(self, f"fc{i + 1}", nn.Linear(dims[i], dims[i + 1]))
            if batch_norm:
                setattr(self, f"bn{i + 1}", nn.BatchNorm1d(dims[i + 1]))
        setattr(self, f"fc{len(hidden) + 1}", nn.Linear(dims[-1], out_features))

    def forward(self, x):
        for i in range(len(self.hidden)):
            x = F.relu(getattr(self, f"fc{i + 1}")(x))
            if self.batch_norm:
                x = getattr(self, f"bn{i + 1}")(x)
        return getattr(self, f"fc{len(self.hidden) + 1}")(x)

    def fused_spec(self):
        from raydp_b200.ops.fused_mlp import MLPSpec
        bn = getattr(self, "bn1", None)
        return MLPSpec(self.in_features, self.hidden, self.out_features, self.batch_norm,
                       bn_eps=bn.eps if bn is not None else 1e-5,
                       bn_momentum=bn.momentum if bn is not None else 0.1)


class NYCTaxiMLP(TabularMLP):
    """The reference's NYC-taxi fare model: 20 engineered features -> fare."""

    def __init__(self, cols: int = 20):
        super().__init__(cols, (256, 128, 64, 16), 1, True)
