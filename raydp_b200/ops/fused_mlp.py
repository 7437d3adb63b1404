"""Whole-step fused trainer for ``Linear -> ReLU -> BatchNorm1d`` chains with a scalar head.

This is the B200-native restatement of the reference's flagship model and loop: the NYC-taxi MLP
(reference ``examples/pytorch_nyctaxi.py:69-92``: 20 -> 256 -> 128 -> 64 -> 16 -> 1, ReLU *then*
BatchNorm after every hidden layer, SmoothL1Loss, Adam) driven by
``TorchEstimator.train_epoch`` (reference ``torch/estimator.py:272-293``).

One training step is ~25 kernel launches and no host synchronisation:

* forward: every hidden layer is one tcgen05 GEMM whose epilogue adds the bias, applies ReLU,
  stores bf16 and accumulates the batch statistics; BatchNorm is folded into the next layer's
  weights by a tiny kernel (``rdp_bn_fold``), so normalised activations are never written;
* head: the last Linear, the loss, its gradient and the metric sums in one pass (``rdp_mlp_head``);
* backward: per layer one MN-major split-K tcgen05 wgrad and one tcgen05 dgrad whose epilogue applies
  the BatchNorm + ReLU backward of the layer it feeds -- the batch sums BN backward needs are derived from
  the weight gradient by an O(N K) kernel (``rdp_bn_bwd_coef``), so no elementwise pass over the
  ``[B, width]`` gradients remains (only the narrow last hidden layer keeps one);
* update: Adam / SGD over ONE flat fp32 parameter buffer (or the fused all-reduce + optimizer
  kernel of :mod:`raydp_b200.parallel` when world_size > 1).

Parameters live in flat fp32 buffers laid out ``[W1, b1, gamma1, beta1, W2, ...]``; views of those
buffers are exported with the same names as the reference model's ``state_dict``
(``fc1.weight`` ... ``bn4.num_batches_tracked``) so the ``model.pt`` checkpoint is loadable by
the user's plain ``nn.Module``.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Dict, List, Sequence, Tuple

import torch

from raydp_b200 import ops


def _round_up(v: int, m: int) -> int:
    return (v + m - 1) // m * m


@dataclass
class MLPSpec:
    in_features: int
    hidden: Sequence[int]
    out_features: int = 1
    batch_norm: bool = True
    bn_eps: float = 1e-5
    bn_momentum: float = 0.1
    linear_names: Sequence[str] = field(default_factory=list)   # state_dict prefixes
    bn_names: Sequence[str] = field(default_factory=list)

    def __post_init__(self):
        n = len(self.hidden)
        if not self.linear_names:
            self.linear_names = [f"fc{i + 1}" for i in range(n + 1)]
        if not self.bn_names:
            self.bn_names = [f"bn{i + 1}" for i in range(n)]
        if self.out_features != 1:
            raise ValueError("the fused head supports one output unit")
        if self.hidden[-1] not in (16, 32, 64):
            raise ValueError("the last hidden width must be 16, 32 or 64 for the fused head")
        for h in self.hidden:
            if h % 16 != 0 or (256 % (h // 8) != 0 and h // 8 <= 256):
                raise ValueError(f"hidden width {h} is not supported by the fused kernels")


class FusedMLP:
    """Device-resident parameters, workspaces and the step functions."""

    def __init__(self, spec: MLPSpec, max_batch: int, device: torch.device, loss: str = "smooth_l1"):
        self.spec = spec
        self.device = device
        self.loss_kind = ops.LOSS_KINDS[loss]
        self.max_batch = int(max_batch)
        # Input rows are padded to a multiple of 32 bf16 = 64 bytes: TMA needs a 16-byte pitch, but a 48-byte
        # pitch (20 features -> 24) made every operand box a bundle of misaligned, mostly out-of-bounds row
        # fetches and the first-layer GEMM ran at half speed (48 us vs 23 us for [131072, 24|32] x [256, .],
        # profiles/gemm_tile_timeline.md)
        self.kpad = _round_up(max(spec.in_features, 8), 32)
        dims = [spec.in_features] + list(spec.hidden)
        self.dims = dims
        L = len(spec.hidden)
        self.L = L
        # ---- flat parameter layout -------------------------------------------------------
        self._slices: Dict[str, Tuple[int, Tuple[int, ...]]] = {}
        off = 0

        def take(name, shape):
            nonlocal off
            n = 1
            for s in shape:
                n *= s
            self._slices[name] = (off, tuple(shape))
            off += n
        for l in range(L):
            take(f"W{l}", (dims[l + 1], dims[l]))
            take(f"b{l}", (dims[l + 1],))
            if spec.batch_norm:
                take(f"g{l}", (dims[l + 1],))
                take(f"be{l}", (dims[l + 1],))
        take("Wh", (1, dims[-1]))
        take("bh", (1,))
        self.n_params = _round_up(off, 4)      # the all-reduce kernel moves 16-byte vectors
        f32 = dict(dtype=torch.float32, device=device)
        bf16 = dict(dtype=torch.bfloat16, device=device)
        self.params = torch.zeros(self.n_params, **f32)
        # one zero-able region: [grads | stats1 | stats2]
        stat_elems = sum(2 * h for h in spec.hidden)
        self._zero_region = torch.zeros(self.n_params + 2 * stat_elems, **f32)
        self.grads = self._zero_region[:self.n_params]
        s1 = self._zero_region[self.n_params:self.n_params + stat_elems]
        s2 = self._zero_region[self.n_params + stat_elems:]
        self.stats1, self.stats2 = [], []
        o = 0
        for h in spec.hidden:
            self.stats1.append(s1[o:o + 2 * h].view(2, h))
            self.stats2.append(s2[o:o + 2 * h].view(2, h))
            o += 2 * h
        self.running_mean = [torch.zeros(h, **f32) for h in spec.hidden]
        self.running_var = [torch.ones(h, **f32) for h in spec.hidden]
        self.num_batches_tracked = 0
        self.bn_a = [torch.ones(h, **f32) for h in spec.hidden]
        self.bn_c = [torch.zeros(h, **f32) for h in spec.hidden]
        self.bn_mean = [torch.zeros(h, **f32) for h in spec.hidden]
        self.bn_rstd = [torch.ones(h, **f32) for h in spec.hidden]
        # bf16 operand copies: W0 (cast), folded W_l (l >= 1), transposed unfolded W_l for dgrad
        self.w_bf = [torch.zeros(dims[1], self.kpad, **bf16)]
        self.w_t = [None]
        self.b_folded = [None]
        for l in range(1, L):
            self.w_bf.append(torch.zeros(dims[l + 1], dims[l], **bf16))
            self.w_t.append(torch.zeros(dims[l], dims[l + 1], **bf16))
            self.b_folded.append(torch.zeros(dims[l + 1], **f32))
        self.head_fold_scratch = torch.zeros(1, dims[-1], **bf16)
        self.head_b_scratch = torch.zeros(1, **f32)
        # activations and gradients
        B = self.max_batch
        self.z = [torch.empty(B, h, **bf16) for h in spec.hidden]
        self.dz = [torch.empty(B, h, **bf16) for h in spec.hidden]
        self.pred = torch.empty(B, **f32)
        self.accum = torch.zeros(8, **f32)
        self._ones = [torch.ones(h, **f32) for h in spec.hidden]
        # per-column constants of the BN+ReLU backward fused into the dgrad epilogue (ops.bn_bwd_coef)
        self.bwd_alpha = [torch.zeros(h, **f32) for h in spec.hidden]
        self.bwd_beta = [torch.zeros(h, **f32) for h in spec.hidden]
        self.bwd_delta = [torch.zeros(h, **f32) for h in spec.hidden]
        import os
        self.fuse_bn_bwd = os.environ.get("RAYDP_B200_FUSE_BN_BWD", "1") != "0"

    # ------------------------------------------------------------------ parameter views
    def view(self, name: str, grad: bool = False) -> torch.Tensor:
        off, shape = self._slices[name]
        n = 1
        for s in shape:
            n *= s
        return (self.grads if grad else self.params)[off:off + n].view(shape)

    def load_state_dict(self, state: Dict[str, torch.Tensor]) -> None:
        sp = self.spec
        with torch.no_grad():
            for l in range(self.L):
                ln = sp.linear_names[l]
                self.view(f"W{l}").copy_(state[f"{ln}.weight"])
                self.view(f"b{l}").copy_(state[f"{ln}.bias"])
                if sp.batch_norm:
                    bn = sp.bn_names[l]
                    self.view(f"g{l}").copy_(state[f"{bn}.weight"])
                    self.view(f"be{l}").copy_(state[f"{bn}.bias"])
                    self.running_mean[l].copy_(state[f"{bn}.running_mean"])
                    self.running_var[l].copy_(state[f"{bn}.running_var"])
                    if f"{bn}.num_batches_tracked" in state:
                        self.num_batches_tracked = int(state[f"{bn}.num_batches_tracked"])
            hn = sp.linear_names[self.L]
            self.view("Wh").copy_(state[f"{hn}.weight"].reshape(1, -1))
            self.view("bh").copy_(state[f"{hn}.bias"].reshape(1))

    def state_dict(self) -> Dict[str, torch.Tensor]:
        sp, out = self.spec, {}
        for l in range(self.L):
            ln = sp.linear_names[l]
            out[f"{ln}.weight"] = self.view(f"W{l}").detach().cpu().clone()
            out[f"{ln}.bias"] = self.view(f"b{l}").detach().cpu().clone()
        hn = sp.linear_names[self.L]
        out[f"{hn}.weight"] = self.view("Wh").detach().cpu().clone()
        out[f"{hn}.bias"] = self.view("bh").detach().cpu().clone()
        if sp.batch_norm:
            for l in range(self.L):
                bn = sp.bn_names[l]
                out[f"{bn}.weight"] = self.view(f"g{l}").detach().cpu().clone()
                out[f"{bn}.bias"] = self.view(f"be{l}").detach().cpu().clone()
                out[f"{bn}.running_mean"] = self.running_mean[l].detach().cpu().clone()
                out[f"{bn}.running_var"] = self.running_var[l].detach().cpu().clone()
                out[f"{bn}.num_batches_tracked"] = torch.tensor(self.num_batches_tracked)
        return out

    # ------------------------------------------------------------------ forward / backward
    def _gamma(self, l):
        return self.view(f"g{l}") if self.spec.batch_norm else self._ones[l]

    def _beta(self, l):
        return self.view(f"be{l}") if self.spec.batch_norm else self.bn_c[l].zero_()

    def _forward(self, x: torch.Tensor, B: int, train: bool) -> List[torch.Tensor]:
        sp = self.spec
        zs = []
        ops.cast_weight(self.view("W0"), self.w_bf[0])
        z = self.z[0][:B]
        ops.gemm_tn(x, self.w_bf[0], z, bias=self.view("b0"), relu=True, k=self.kpad,
                    stats_mode=1 if (train and sp.batch_norm) else 0, stats=self.stats1[0])
        zs.append(z)
        for l in range(1, self.L):
            self._fold(l - 1, B, train, self.view(f"W{l}"), self.view(f"b{l}"), self.w_bf[l], self.w_t[l],
                       self.b_folded[l])
            z = self.z[l][:B]
            ops.gemm_tn(zs[-1], self.w_bf[l], z, bias=self.b_folded[l], relu=True,
                        stats_mode=1 if (train and sp.batch_norm) else 0, stats=self.stats1[l])
            zs.append(z)
        self._fold(self.L - 1, B, train, self.view("Wh"), self.view("bh"), self.head_fold_scratch, None,
                   self.head_b_scratch)
        return zs

    def _fold(self, l, B, train, w, b, w_folded, w_t, b_folded) -> None:
        sp = self.spec
        if sp.batch_norm:
            ops.bn_fold(self.stats1[l], B, self._gamma(l), self._beta(l), self.running_mean[l],
                        self.running_var[l], sp.bn_momentum, sp.bn_eps, not train, self.bn_a[l], self.bn_c[l],
                        self.bn_mean[l], self.bn_rstd[l], w, b, w_folded, w_t, b_folded)
        else:
            ops.cast_weight(w, w_folded, w_t)
            b_folded.copy_(b)

    def group_ranges(self) -> List[Tuple[int, int]]:
        """``[lo, hi)`` of every layer's parameters in the flat buffers, layer 0 first; the last range also holds
        the head.  Gradients of range ``l`` are final when ``train_step`` calls ``on_group_ready(l)``."""
        starts = [self._slices[f"W{l}"][0] for l in range(self.L)]
        return [(starts[l], starts[l + 1] if l + 1 < self.L else self.n_params) for l in range(self.L)]

    def train_step(self, x: torch.Tensor, target: torch.Tensor, on_group_ready=None) -> None:
        """Forward + backward for one batch; gradients land in ``self.grads`` (sum over the local
        batch of d(mean loss)); loss / metric sums are added to ``self.accum``.  ``on_group_ready(l)`` is
        called as soon as every gradient of layer ``l`` (see :meth:`group_ranges`) is final -- last layer
        first -- so a data-parallel caller can reduce and update that layer while backward goes on."""
        B = x.shape[0]
        assert B <= self.max_batch and x.dtype == torch.bfloat16 and x.shape[1] == self.kpad
        sp = self.spec
        self._zero_region.zero_()
        zs = self._forward(x, B, True)
        L = self.L
        bn = sp.batch_norm
        dz = [d[:B] for d in self.dz]
        ops.mlp_head(zs[-1], self.bn_a[L - 1] if bn else None, self.bn_c[L - 1] if bn else None,
                     self.view("Wh"), self.view("bh"), target, self.loss_kind, 1.0 / B, self.pred[:B],
                     self.accum, dz[L - 1], self.stats2[L - 1], self.view("Wh", True), self.view("bh", True))
        # Backward.  dz[l] arrives either as dL/dy_l (gradient w.r.t. the BN output: from the head, or from a
        # plain dgrad) and is turned into dL/dpre_l by the elementwise BN+ReLU backward -- or, for layers the
        # fused path covers, ALREADY as dL/dpre_l: the dgrad GEMM of the layer above applied BN+ReLU backward in
        # its epilogue, with batch statistics derived from that layer's weight gradient (ops.bn_bwd_coef).
        pre_ready = False
        for l in range(L - 1, -1, -1):
            if not pre_ready:
                ops.bn_relu_bwd(dz[l], zs[l], dz[l], self.stats2[l], self.bn_mean[l], self.bn_rstd[l],
                                self.bn_a[l], self.view(f"g{l}", True) if bn else self.bn_a[l],
                                self.view(f"be{l}", True) if bn else self.bn_c[l], self.view(f"b{l}", True), bn)
            inp = zs[l - 1] if l > 0 else x
            ops.gemm_wgrad(dz[l], inp, self.view(f"W{l}", True), nw=self.dims[l])
            pre_ready = False
            if l > 0:
                width = self.dims[l]
                if bn and self.fuse_bn_bwd and width % 64 == 0 and width <= 256:
                    ops.bn_bwd_coef(self.view(f"W{l}", True), self.view(f"b{l}", True), self.w_t[l],
                                    self.bn_a[l - 1], self.bn_c[l - 1], self.bn_mean[l - 1], self.bn_rstd[l - 1], B,
                                    self.view(f"g{l - 1}", True), self.view(f"be{l - 1}", True),
                                    self.bwd_alpha[l - 1], self.bwd_beta[l - 1], self.bwd_delta[l - 1])
                    if on_group_ready is not None:
                        on_group_ready(l)
                    ops.gemm_tn_bnbwd(dz[l], self.w_t[l], dz[l - 1], zs[l - 1], self.bwd_alpha[l - 1],
                                      self.bwd_beta[l - 1], self.bwd_delta[l - 1], self.view(f"b{l - 1}", True))
                    pre_ready = True
                    continue
                if bn:
                    ops.grad_fixup(self.view(f"W{l}", True), self.view(f"b{l}", True), self.bn_a[l - 1],
                                   self.bn_c[l - 1])
                if on_group_ready is not None:
                    on_group_ready(l)
                ops.gemm_tn(dz[l], self.w_t[l], dz[l - 1], aux=zs[l - 1],
                            stats_mode=2 if bn else 0, stats=self.stats2[l - 1],
                            aux_relu_mask=False)
            elif on_group_ready is not None:
                on_group_ready(0)
        if bn:
            self.num_batches_tracked += 1

    def eval_step(self, x: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        B = x.shape[0]
        zs = self._forward(x, B, False)
        bn = self.spec.batch_norm
        L = self.L
        ops.mlp_head(zs[-1], self.bn_a[L - 1] if bn else None, self.bn_c[L - 1] if bn else None,
                     self.view("Wh"), self.view("bh"), target, self.loss_kind, 0.0, self.pred[:B],
                     self.accum, None, None, None, None)
        return self.pred[:B]

    def predict(self, x: torch.Tensor) -> torch.Tensor:
        dummy = torch.zeros(x.shape[0], device=x.device)
        keep = self.accum.clone()
        out = self.eval_step(x, dummy).clone()
        self.accum.copy_(keep)
        return out

    def pack_input(self, x: torch.Tensor) -> torch.Tensor:
        """fp32/bf16 ``[B, in_features]`` -> zero-padded bf16 ``[B, kpad]`` (tests and SPMD callers
        that already hold device tensors; the estimator path gets this layout from ingest)."""
        out = torch.zeros(x.shape[0], self.kpad, dtype=torch.bfloat16, device=self.device)
        out[:, :x.shape[1]] = x.to(torch.bfloat16)
        return out
