"""``FusedDataParallel``: data parallelism for arbitrary modules with the all-reduce fused into the
optimizer and overlapped with backward.

Replaces torch DDP + ``optimizer.step()`` as the reference gets them through
``train.torch.prepare_model`` (reference ``torch/estimator.py:243,284-286``) for models that are not on
the fully fused MLP path.

* Parameters are re-homed into flat fp32 buckets (reverse registration order = backward order);
  gradients are views of **registered symmetric memory**, so backward writes them where the peers
  read them -- no bucket copy.
* When the last gradient of a bucket lands, the bucket's kernel (NVLS ``multimem`` reduction inside
  the NVSwitch where available, else two-shot reduce-scatter by pull + all-gather by push over
  NVLink; then Adam/SGD on the reduced gradient and the bf16 operand shadow, one launch) is
  enqueued on a communication stream, behind an event from the compute stream.  The kernel is
  launched in a 256-thread / 40-register shape so its blocks are co-resident with the persistent
  tcgen05 GEMM CTAs: buckets of late layers reduce and update while earlier layers are still in backward.
* ``world_size == 1`` degenerates to the fused flat optimizer.
"""
from __future__ import annotations

import os
from typing import Dict, List, Optional

import torch
import torch.distributed as dist
import torch.nn as nn

from raydp_b200 import ops


class _Bucket:
    def __init__(self, params: List[nn.Parameter], device: torch.device, world: int, grad_wire: str = "fp32"):
        self.params = params
        self.offsets, off = [], 0
        for p in params:
            self.offsets.append(off)
            off += (p.numel() + 7) // 8 * 8      # 16-byte aligned in the bf16 shadow too (TMA base)
        self.n = off
        f32 = dict(dtype=torch.float32, device=device)
        self.flat = torch.zeros(self.n, **f32)
        self.m = torch.zeros(self.n, **f32)
        self.v = torch.zeros(self.n, **f32)
        # bf16 copy of the parameters, rewritten by the optimizer kernel itself: FusedLinear layers
        # take their GEMM operands from it, so no fp32 -> bf16 cast pass runs per step
        self.shadow = torch.zeros(self.n, dtype=torch.bfloat16, device=device)
        self.reducer = None
        if world > 1:
            from raydp_b200.comm import NvlsReducer, P2PReducer, nvls_available
            # Measured (profiles/README.md): the in-switch reduction wins from 4 ranks up (8 GPUs, 1 GB:
            # 2.28 ms vs 3.20 ms two-shot); with 2 ranks the peer-memory two-shot is faster (1.69 vs 2.7 ms).
            use_nvls = world >= 4 and nvls_available(device) and os.environ.get("RAYDP_B200_NO_NVLS") != "1"
            if grad_wire == "bf16":
                # bf16 on the wire, fp32 accumulate: gradients stay in a local fp32 buffer and are rounded once
                # into the (half-size) symmetric buffer by the reduction kernel itself
                if use_nvls:
                    self.reducer = NvlsReducer(self.n, device, wire="bf16")
                else:
                    self.reducer = P2PReducer(torch.zeros(self.n, **f32), device, algo="bf16")
            elif use_nvls:
                # reduction inside the NVSwitch (multimem): gradients live in the multicast-mapped buffer
                self.reducer = NvlsReducer(self.n, device)
            else:
                self.reducer = P2PReducer(None, device, algo="two_shot_fixed", n=self.n)
            self.grads = self.reducer.grads
        else:
            self.grads = torch.zeros(self.n, **f32)
        with torch.no_grad():
            for p, o in zip(params, self.offsets):
                self.flat[o:o + p.numel()].copy_(p.detach().reshape(-1))
                p.data = self.flat[o:o + p.numel()].view(p.shape)
                p.grad = self.grads[o:o + p.numel()].view(p.shape)
            self.shadow.copy_(self.flat)
        self.pending = len(params)
        self.done_event: Optional[torch.cuda.Event] = None

    def offset_of(self, p: nn.Parameter) -> int:
        return self.offsets[[id(q) for q in self.params].index(id(p))]

    def shadow_of(self, p: nn.Parameter) -> torch.Tensor:
        o = self.offset_of(p)
        return self.shadow[o:o + p.numel()].view(p.shape)


class FusedDataParallel(nn.Module):
    def __init__(self, module: nn.Module, optimizer: torch.optim.Optimizer, device: Optional[torch.device] = None,
                 bucket_mb: int = 512, overlap: bool = True, grad_wire: Optional[str] = None):
        """``grad_wire``: ``"fp32"`` (default, what DDP's buckets carry) or ``"bf16"`` -- gradients cross NVLink as
        bf16 and are summed in fp32 (half the link bytes; the counterpart of DDP's bf16 compression hook).
        Environment default: ``RAYDP_B200_GRAD_WIRE``."""
        super().__init__()
        self.grad_wire = grad_wire or os.environ.get("RAYDP_B200_GRAD_WIRE", "fp32")
        assert self.grad_wire in ("fp32", "bf16"), self.grad_wire
        self.module = module
        self.device = device or next(module.parameters()).device
        self.world = dist.get_world_size() if dist.is_initialized() else 1
        self.overlap = overlap
        g = optimizer.param_groups[0]
        plain = len(optimizer.param_groups) == 1 and not g.get("maximize") and not g.get("amsgrad")
        if plain and type(optimizer) in (torch.optim.Adam, torch.optim.AdamW):
            self.kind = 1
            self.hp = dict(betas=g["betas"], eps=g["eps"], weight_decay=g["weight_decay"],
                           adamw=type(optimizer) is torch.optim.AdamW)
        elif plain and type(optimizer) is torch.optim.SGD:
            self.kind = 2
            self.hp = dict(weight_decay=g["weight_decay"], momentum=g["momentum"], dampening=g["dampening"],
                           nesterov=g["nesterov"])
        else:
            # any other torch.optim.Optimizer (the reference accepts all of them, torch/estimator.py:190-202):
            # the bucket kernels only average the gradients over the ranks and the user's own
            # optimizer.step() runs afterwards on the re-homed parameters
            self.kind = 0
            self.hp = {}
        self.optimizer = optimizer
        params = [p for p in module.parameters() if p.requires_grad]
        if self.world > 1:
            for p in params:
                dist.broadcast(p.data, src=0)
        limit = bucket_mb * (1 << 20) // 4
        self.buckets: List[_Bucket] = []
        cur, size = [], 0
        for p in reversed(params):
            if cur and size + p.numel() > limit:
                self.buckets.append(_Bucket(cur, self.device, self.world, self.grad_wire))
                cur, size = [], 0
            cur.append(p)
            size += p.numel()
        if cur:
            self.buckets.append(_Bucket(cur, self.device, self.world, self.grad_wire))
        self._bucket_of: Dict[int, _Bucket] = {id(p): b for b in self.buckets for p in b.params}
        f32 = dict(dtype=torch.float32, device=self.device)
        self.step_dev = torch.zeros(1, **f32)
        self.lr_dev = torch.zeros(1, **f32)
        self.comm_stream = torch.cuda.Stream(device=self.device)
        if self.world > 1 and overlap:
            ops.set_dynamic_tiles(True)     # backward GEMMs share the SMs with the bucket all-reduces
        from raydp_b200.ops.linear import FusedLinear
        self._fused_linears = [m for m in module.modules() if isinstance(m, FusedLinear)]
        fused_w = {id(m.weight) for m in self._fused_linears}
        fused_b = {id(m.bias) for m in self._fused_linears if m.bias is not None}
        for m in self._fused_linears:
            m._grad_ready_hooks.append(self._fused_linear_ready)
            if self.kind != 0 and not m.padded:
                m.use_external_shadow(self._bucket_of[id(m.weight)].shadow_of(m.weight))
        self._seed_state(params)
        for p in params:
            if id(p) not in fused_w and id(p) not in fused_b:
                p.register_post_accumulate_grad_hook(self._param_ready)

    def _seed_state(self, params) -> None:
        """An optimizer instance may arrive with state (Adam moments of an earlier fit): carry it over."""
        if self.kind == 0 or not self.optimizer.state:
            return
        step = 0.0
        with torch.no_grad():
            for p in params:
                st = self.optimizer.state.get(p)
                if not st:
                    continue
                b = self._bucket_of[id(p)]
                o, n = b.offset_of(p), p.numel()
                if "exp_avg" in st:
                    b.m[o:o + n].copy_(st["exp_avg"].reshape(-1))
                    b.v[o:o + n].copy_(st["exp_avg_sq"].reshape(-1))
                elif st.get("momentum_buffer") is not None:
                    b.m[o:o + n].copy_(st["momentum_buffer"].reshape(-1))
                    step = max(step, 1.0)
                if "step" in st:
                    step = max(step, float(st["step"]))
        if step:
            self.step_dev.fill_(step)

    # ------------------------------------------------------------------ hooks
    def _fused_linear_ready(self, layer) -> None:
        self._param_ready(layer.weight)
        if layer.bias is not None:
            self._param_ready(layer.bias)

    def _param_ready(self, p) -> None:
        b = self._bucket_of[id(p)]
        b.pending -= 1
        if b.pending == 0:
            self._reduce_and_update(b)

    def _opt_args(self, b: _Bucket):
        from raydp_b200.comm import make_opt_args
        if self.kind == 1:
            return make_opt_args(1, b.flat, b.m, b.v, lr_dev=self.lr_dev, step_dev=self.step_dev, shadow=b.shadow,
                                 **self.hp)
        return make_opt_args(2, b.flat, b.m if self.hp["momentum"] else None, None, lr_dev=self.lr_dev,
                             step_dev=self.step_dev, shadow=b.shadow, **self.hp)

    def _reduce_and_update(self, b: _Bucket) -> None:
        compute = torch.cuda.current_stream(self.device)
        stream = self.comm_stream if (self.overlap and self.world > 1) else compute
        if stream is not compute:
            ready = torch.cuda.Event()
            ready.record(compute)
            stream.wait_event(ready)
        with torch.cuda.stream(stream):
            if self.kind == 0:
                if b.reducer is not None:
                    b.reducer.allreduce_(1.0 / self.world)
            elif b.reducer is not None:
                if getattr(b.reducer, "kind", "") == "nvls":
                    b.reducer.allreduce_step_(self._opt_args(b), 1.0 / self.world, overlap=stream is not compute)
                else:
                    b.reducer.allreduce_step_(self._opt_args(b), 1.0 / self.world)
            elif self.kind == 1:
                ops.adam_step(b.flat, b.grads, b.m, b.v, 0.0, self.hp["betas"][0], self.hp["betas"][1], self.hp["eps"],
                              self.hp["weight_decay"], 1, 1.0, self.hp["adamw"], self.lr_dev, self.step_dev,
                              shadow=b.shadow)
            else:
                ops.sgd_step(b.flat, b.grads, b.m if self.hp["momentum"] else None, 0.0, self.hp["momentum"],
                             self.hp["dampening"], self.hp["weight_decay"], self.hp["nesterov"], False, 1.0,
                             self.lr_dev, self.step_dev, shadow=b.shadow)
            if self.kind != 0:
                b.grads.zero_()
            b.done_event = torch.cuda.Event()
            b.done_event.record(stream)

    # ------------------------------------------------------------------ step API
    def forward(self, *args, **kwargs):
        return self.module(*args, **kwargs)

    def backward_and_step(self, loss: torch.Tensor) -> None:
        """``loss.backward()`` with bucket all-reduce + optimizer fused in; returns with every update
        ordered before later work on the current stream."""
        self.lr_dev.fill_(float(self.optimizer.param_groups[0]["lr"]))
        ops.add_scalar(self.step_dev, 1.0)
        for b in self.buckets:
            b.pending = len(b.params)
        loss.backward()
        compute = torch.cuda.current_stream(self.device)
        for b in self.buckets:
            if b.pending != 0:                      # parameters that got no gradient this step
                b.pending = 0
                self._reduce_and_update(b)
            if b.done_event is not None:
                compute.wait_event(b.done_event)
        if self.kind == 0:
            self.optimizer.step()                   # gradients are the rank average by now
            for b in self.buckets:
                b.grads.zero_()
        for m in self._fused_linears:
            m.invalidate_shadow()                   # parameters were updated by our kernel, not by torch

    def state_dict(self, *args, **kwargs):
        return self.module.state_dict(*args, **kwargs)

    def close(self) -> None:
        for b in self.buckets:
            if b.reducer is not None:
                b.reducer.close()
