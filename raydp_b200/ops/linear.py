"""``FusedLinear``: ``nn.Linear`` whose forward / dgrad / wgrad run on the tcgen05 GEMM kernels.

Generic building block for user models outside the fully fused MLP path (SURVEY.md section 8 K6:
"wire as custom autograd ops for nn.Linear-shaped layers").  Parameters stay fp32 (master copy,
same ``weight`` / ``bias`` names as ``nn.Linear`` so state_dicts are interchangeable); a bf16
shadow of the weight (and its transpose for dgrad) is refreshed when the parameter changes.
The weight gradient is accumulated in fp32 *directly into* ``weight.grad`` by the split-K MN-major
kernel, which is what lets :class:`raydp_b200.parallel.FusedDataParallel` place gradients in
registered symmetric memory without a copy.
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from raydp_b200 import ops


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, layer, relu):
        x2 = x.reshape(-1, x.shape[-1])
        if x2.dtype != torch.bfloat16:
            x2 = x2.to(torch.bfloat16)
        x2 = x2.contiguous()
        w_bf, _ = layer._shadows()
        out = torch.empty(x2.shape[0], layer.out_features, dtype=torch.bfloat16, device=x.device)
        ops.gemm_tn(x2, w_bf, out, bias=bias, relu=relu)
        ctx.layer, ctx.relu = layer, relu
        ctx.save_for_backward(x2, out if relu else None)
        ctx.in_shape = x.shape
        return out.reshape(*x.shape[:-1], layer.out_features)

    @staticmethod
    def backward(ctx, grad_out):
        layer = ctx.layer
        x2, out = ctx.saved_tensors
        g = grad_out.reshape(-1, layer.out_features)
        if g.dtype != torch.bfloat16:
            g = g.to(torch.bfloat16)
        g = g.contiguous()
        if ctx.relu:
            g = torch.where(out > 0, g, torch.zeros_like(g))
        _, w_t = layer._shadows()
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(x2.shape[0], layer.in_features, dtype=torch.bfloat16, device=g.device)
            ops.gemm_tn(g, w_t, dx)
            dx = dx.reshape(ctx.in_shape)
        # fp32 accumulation straight into the parameter's gradient buffer
        w = layer.weight
        if w.grad is None:
            w.grad = torch.zeros_like(w)
        ops.gemm_wgrad(g, x2, w.grad)
        if layer.bias is not None:
            db = g.float().sum(0)
            if layer.bias.grad is None:
                layer.bias.grad = db
            else:
                layer.bias.grad.add_(db)
        layer._fire_grad_hooks()
        return dx, None, None, None, None


class FusedLinear(nn.Module):
    """Drop-in ``nn.Linear`` (optionally fused with ReLU) on sm_100a tensor cores.
    ``in_features`` must be a multiple of 8 and ``out_features`` a multiple of 16."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, relu: bool = False):
        super().__init__()
        if in_features % 8 or out_features % 16:
            raise ValueError("FusedLinear needs in_features % 8 == 0 and out_features % 16 == 0")
        self.in_features, self.out_features, self.relu = in_features, out_features, relu
        ref = nn.Linear(in_features, out_features, bias=bias)
        self.weight = nn.Parameter(ref.weight.detach().clone())
        self.bias = nn.Parameter(ref.bias.detach().clone()) if bias else None
        self._w_bf: Optional[torch.Tensor] = None
        self._w_t: Optional[torch.Tensor] = None
        self._shadow_version = -1
        self._grad_ready_hooks = []

    def _shadows(self):
        w = self.weight
        if self._w_bf is None or self._w_bf.device != w.device:
            self._w_bf = torch.empty(self.out_features, self.in_features, dtype=torch.bfloat16, device=w.device)
            self._w_t = torch.empty(self.in_features, self.out_features, dtype=torch.bfloat16, device=w.device)
            self._shadow_version = -1
        if self._shadow_version != w._version:
            ops.cast_weight(w.detach(), self._w_bf, self._w_t)
            self._shadow_version = w._version
        return self._w_bf, self._w_t

    def invalidate_shadow(self) -> None:
        """Call after the parameter storage was updated by a kernel torch does not see."""
        self._shadow_version = -1

    def _fire_grad_hooks(self) -> None:
        for hook in self._grad_ready_hooks:
            hook(self)

    def forward(self, x):
        if not x.is_cuda:
            y = torch.nn.functional.linear(x, self.weight, self.bias)
            return torch.relu(y) if self.relu else y
        return _LinearFn.apply(x, self.weight, self.bias, self, self.relu)
