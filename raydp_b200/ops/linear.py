"""``FusedLinear``: ``nn.Linear`` whose forward / dgrad / wgrad run on the tcgen05 GEMM kernels.

Generic building block for user models outside the fully fused MLP path (SURVEY.md section 8 K6:
"wire as custom autograd ops for nn.Linear-shaped layers").  Parameters stay fp32 (master copy,
same ``weight`` / ``bias`` names as ``nn.Linear`` so state_dicts are interchangeable); a bf16
shadow of the weight (and its transpose for dgrad) is refreshed when the parameter changes.
The weight gradient is accumulated in fp32 *directly into* ``weight.grad`` by the split-K MN-major
kernel, which is what lets :class:`raydp_b200.parallel.FusedDataParallel` place gradients in
registered symmetric memory without a copy.

Any ``in_features`` / ``out_features`` are accepted: TMA wants a 16-byte row pitch and the UMMA tile a
multiple of 16 output columns, so odd shapes (a 20-feature input layer, a 1-unit head) run with the
bf16 operand copies zero-padded to ``[ceil16(out), ceil16(in)]``; the fp32 parameters and gradients
keep their exact shapes.  ``FusedLinear.from_linear`` adopts an existing ``nn.Linear``'s Parameter
objects, which is how the generic GPU engine puts an arbitrary user module on these kernels
(reference ``torch/estimator.py:180-187`` accepts any module).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.nn as nn

from raydp_b200 import ops


_PAIR_MIN_OUT = 37 * 256 * 256      # outputs below this stay on the 1-CTA kernels


def _up(v: int, m: int) -> int:
    return (v + m - 1) // m * m


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, layer, relu):
        K, N, Kp, Np = layer.in_features, layer.out_features, layer._kp, layer._np
        x2 = x.reshape(-1, K)
        if Kp != K:
            xp = torch.zeros(x2.shape[0], Kp, dtype=torch.bfloat16, device=x.device)
            xp[:, :K] = x2
            x2 = xp
        else:
            if x2.dtype != torch.bfloat16:
                x2 = x2.to(torch.bfloat16)
            x2 = x2.contiguous()
        w_bf, _ = layer._shadows(need_transpose=False)
        out = torch.empty(x2.shape[0], Np, dtype=torch.bfloat16, device=x.device)
        ops.gemm_tn(x2, w_bf, out, bias=layer._bias_operand(), relu=relu)
        ctx.layer, ctx.relu = layer, relu
        ctx.save_for_backward(x2, out if relu else None)
        ctx.in_shape = x.shape
        y = out if Np == N else out[:, :N]
        return y.reshape(*x.shape[:-1], N)

    @staticmethod
    def backward(ctx, grad_out):
        layer = ctx.layer
        K, N, Kp, Np = layer.in_features, layer.out_features, layer._kp, layer._np
        x2, out = ctx.saved_tensors
        g = grad_out.reshape(-1, N)
        if Np != N:
            gp = torch.zeros(g.shape[0], Np, dtype=torch.bfloat16, device=g.device)
            gp[:, :N] = g
            g = gp
        else:
            if g.dtype != torch.bfloat16:
                g = g.to(torch.bfloat16)
            g = g.contiguous()
        # one pass: ReLU mask (when the layer is fused with ReLU) + bias gradient column sums
        db = None
        if layer.bias is not None:
            if layer.bias.grad is None:
                layer.bias.grad = torch.zeros_like(layer.bias)
            db = layer.bias.grad if Np == N else torch.zeros(Np, dtype=torch.float32, device=g.device)
        if ctx.relu or db is not None:
            g = ops.relu_bwd_bias(g, out if ctx.relu else None, db)
        if db is not None and Np != N:
            layer.bias.grad.add_(db[:N])
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty(x2.shape[0], Kp, dtype=torch.bfloat16, device=g.device)
            if Kp == K and Np == N and K % 256 == 0 and x2.shape[0] * K >= _PAIR_MIN_OUT:
                # dX = dY . W with the row-major weight as an MN-major operand: no transposed copy
                w_bf, _ = layer._shadows(need_transpose=False)
                ops.gemm2_nn(g, w_bf, dx)
            else:
                _, w_t = layer._shadows(need_transpose=True)
                ops.gemm_tn(g, w_t, dx)
            if Kp != K:
                dx = dx[:, :K]
            dx = dx.reshape(ctx.in_shape)
        # fp32 accumulation straight into the parameter's gradient buffer
        w = layer.weight
        if w.grad is None:
            w.grad = torch.zeros_like(w)
        ops.gemm_wgrad(g, x2, w.grad, mw=N, nw=K)
        layer._fire_grad_hooks()
        return dx, None, None, None, None


class FusedLinear(nn.Module):
    """Drop-in ``nn.Linear`` (optionally fused with ReLU) on sm_100a tensor cores."""

    def __init__(self, in_features: int, out_features: int, bias: bool = True, relu: bool = False,
                 _adopt: Optional[nn.Linear] = None):
        super().__init__()
        self.in_features, self.out_features, self.relu = in_features, out_features, relu
        # the input width is the N of the dgrad GEMM, so it is padded to the UMMA granularity as well
        self._kp, self._np = _up(in_features, 16), _up(out_features, 16)
        if _adopt is not None:
            self.weight = _adopt.weight
            self.bias = _adopt.bias
        else:
            ref = nn.Linear(in_features, out_features, bias=bias)
            self.weight = nn.Parameter(ref.weight.detach().clone())
            self.bias = nn.Parameter(ref.bias.detach().clone()) if bias else None
        self._w_bf: Optional[torch.Tensor] = None
        self._w_t: Optional[torch.Tensor] = None
        self._bias_pad: Optional[torch.Tensor] = None
        self._shadow_version = -1
        self._external: Optional[torch.Tensor] = None
        self._t_version, self._t_stale = -1, True
        self._grad_ready_hooks = []

    @classmethod
    def from_linear(cls, lin: nn.Linear, relu: bool = False) -> "FusedLinear":
        """A FusedLinear that shares ``lin``'s Parameter objects (optimizers and ``state_dict`` names
        built around the original layer stay valid)."""
        return cls(lin.in_features, lin.out_features, lin.bias is not None, relu, _adopt=lin)

    @property
    def padded(self) -> bool:
        return self._kp != self.in_features or self._np != self.out_features

    def _bias_operand(self) -> Optional[torch.Tensor]:
        if self.bias is None:
            return None
        if self._np == self.out_features:
            return self.bias.detach()
        if self._bias_pad is None or self._bias_pad.device != self.bias.device:
            self._bias_pad = torch.zeros(self._np, dtype=torch.float32, device=self.bias.device)
        self._bias_pad[:self.out_features].copy_(self.bias.detach())
        return self._bias_pad

    def _shadows(self, need_transpose: bool = True):
        """bf16 operand copies of the weight: ``(w_bf [out, in], w_t [in, out] or None)`` (zero-padded to
        the tile granularity for odd shapes).  With an external shadow (see :meth:`use_external_shadow`)
        ``w_bf`` is maintained by the optimizer kernel and never re-cast here; the transposed copy is only
        built for shapes whose dgrad cannot take the row-major weight directly."""
        w = self.weight
        if self._external is not None:
            if need_transpose:
                if self._w_t is None:
                    self._w_t = torch.empty(self.in_features, self.out_features, dtype=torch.bfloat16, device=w.device)
                if self._t_version != w._version or self._t_stale:
                    self._w_t.copy_(self._external.t())
                    self._t_version, self._t_stale = w._version, False
            return self._external, self._w_t
        if self._w_bf is None or self._w_bf.device != w.device:
            self._w_bf = torch.zeros(self._np, self._kp, dtype=torch.bfloat16, device=w.device)
            self._shadow_version = -1
        if need_transpose and (self._w_t is None or self._w_t.device != w.device):
            self._w_t = torch.zeros(self._kp, self._np, dtype=torch.bfloat16, device=w.device)
            self._shadow_version = -1
        if self._shadow_version != w._version or (need_transpose and self._t_stale):
            ops.cast_weight(w.detach(), self._w_bf, self._w_t if need_transpose else None)
            self._shadow_version = w._version
            self._t_stale = not need_transpose
        return self._w_bf, self._w_t

    def use_external_shadow(self, shadow: torch.Tensor) -> None:
        """``shadow``: a bf16 ``[out, in]`` tensor that some other kernel (the fused optimizer) keeps
        equal to the weight; it must already hold the current values.  Unpadded shapes only."""
        assert shadow.shape == self.weight.shape and shadow.dtype == torch.bfloat16 and not self.padded
        self._external = shadow
        self._w_bf = None

    def invalidate_shadow(self) -> None:
        """Call after the parameter storage was updated by a kernel torch does not see."""
        self._shadow_version = -1
        self._t_stale = True

    def _fire_grad_hooks(self) -> None:
        for hook in self._grad_ready_hooks:
            hook(self)

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features}, relu={self.relu}"

    def forward(self, x):
        if not x.is_cuda:
            y = torch.nn.functional.linear(x, self.weight, self.bias)
            return torch.relu(y) if self.relu else y
        return _LinearFn.apply(x, self.weight, self.bias, self, self.relu)


def swap_linears(module: nn.Module) -> int:
    """Replace every plain ``nn.Linear`` under ``module`` by a :class:`FusedLinear` adopting its
    parameters; ``Linear -> ReLU`` neighbours inside an ``nn.Sequential`` become one fused layer (the
    ReLU is applied in the GEMM epilogue and replaced by ``nn.Identity``).  Returns the number swapped."""
    n = 0
    for parent in list(module.modules()):
        names = list(parent._modules.keys())
        for i, name in enumerate(names):
            child = parent._modules[name]
            if type(child) is not nn.Linear:
                continue
            relu = False
            if isinstance(parent, nn.Sequential) and i + 1 < len(names) \
                    and type(parent._modules[names[i + 1]]) is nn.ReLU:
                relu = True
                parent._modules[names[i + 1]] = nn.Identity()
            parent._modules[name] = FusedLinear.from_linear(child, relu=relu)
            n += 1
    return n
