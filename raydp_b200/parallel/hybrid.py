"""Hybrid parallel DLRM: dense layers data-parallel, embedding tables model-parallel.

New capability relative to the reference (the reference keeps all 26 tables
replicated under DDP, ``examples/pytorch_dlrm.ipynb`` cell 9; ``BASELINE.json`` config 3 asks for
table-wise sharding with the all-to-all fused into the lookup).

* Tables are dealt round-robin to ranks.  Every rank publishes its indices transposed ([column, sample]).
  Forward: the owner reads each requesting rank's indices over NVLink with 16-byte loads, gathers from its
  shard, pools multi-hot bags (sum / mean) and stores one vector per (sample, table) straight into that
  rank's feature slab (``rdp_embedding_bag_fwd``); backward: the owner pulls gradient rows from every rank's
  slab, combines duplicate rows in 64-bit fixed point (bit-reproducible) and updates every touched row once
  with SGD, row-wise Adagrad or lazy Adam (``rdp_embedding_bag_bwd``).  The batch of every rank is read from
  that rank's memory, so ranks may hold different (even empty) batches.  Three flag rounds per step delimit
  the phases.
* Bottom / top MLPs run on the tcgen05 GEMM with bias+ReLU epilogues; the backward dgrad GEMM applies
  the ReLU mask in its epilogue and its tensor-core column sums are the bias gradients.
* Interaction: ``rdp_dot_interaction_fwd/bwd``; loss: ``rdp_bce_logits``.
* Dense gradients: peer-memory all-reduce fused with SGD (``raydp_b200.comm``), or a local SGD step.

``LocalWorld`` runs W ranks sequentially on one GPU with all "peer" pointers local: the same
kernels and the same sharding logic, used by the tests where only one GPU exists.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence

import torch

from raydp_b200 import ops
from raydp_b200.models.dlrm import DLRMSpec


def _r8(v: int) -> int:
    return (v + 7) // 8 * 8


class FusedDLRM:
    """One rank's share of the model plus its workspaces."""

    def __init__(self, spec: DLRMSpec, batch: int, device: torch.device, rank: int = 0, world: int = 1,
                 table_optimizer: str = "sgd", table_opt_args: Optional[dict] = None,
                 deterministic: Optional[bool] = None):
        import os
        self.spec, self.B, self.device, self.rank, self.world = spec, int(batch), device, rank, world
        # bit-reproducible steps: weight gradients by ordered split sums, bias gradients by ordered column sums
        # (the table update is always order-independent); RAYDP_B200_DETERMINISTIC=0 selects the atomic epilogues
        self.deterministic = (os.environ.get("RAYDP_B200_DETERMINISTIC", "1") != "0") if deterministic is None \
            else bool(deterministic)
        D, T = spec.embedding_dim, len(spec.table_sizes)
        self.D, self.T, self.S = D, T, T + 1
        self.lens = [int(v) for v in (getattr(spec, "bag_sizes", None) or [1] * T)]
        self.col0 = [sum(self.lens[:t]) for t in range(T)]
        self.C = sum(self.lens)                       # index columns per sample
        self.mean = getattr(spec, "pooling", "sum") == "mean"
        self.ldb = (self.B + 3) // 4 * 4              # pitch of the transposed index matrix
        self.kpad = _r8(spec.num_dense)
        # width of the interaction vector: the module pads it to 8 (models/dlrm.py), the GEMMs want 16
        self.inter_w_model = _r8(self.S * (self.S - 1) // 2 + D + 1)
        self.inter_w = (self.inter_w_model + 15) // 16 * 16
        f32 = dict(dtype=torch.float32, device=device)
        bf = dict(dtype=torch.bfloat16, device=device)
        # ---- sharded tables ------------------------------------------------------------------
        self.owned = [t for t in range(T) if t % world == rank]
        self.tables: Dict[int, torch.Tensor] = {t: torch.zeros(spec.table_sizes[t], D, **f32) for t in self.owned}
        # sparse update: compact fixed-point accumulators (a slot per touched row), row -> slot map, touched list
        if table_optimizer not in ops.EMB_OPT_KINDS:
            raise ValueError(f"table_optimizer must be one of {sorted(ops.EMB_OPT_KINDS)}, got {table_optimizer!r}")
        self.table_opt = ops.EMB_OPT_KINDS[table_optimizer]
        self.table_opt_args = dict(eps=1e-8 if self.table_opt == 2 else 1e-10, beta1=0.9, beta2=0.999,
                                   weight_decay=0.0)
        self.table_opt_args.update(table_opt_args or {})
        i32 = dict(dtype=torch.int32, device=device)
        slots = max(1, sum(world * self.B * self.lens[t] for t in self.owned))      # indices of one step
        self.acc = torch.zeros(slots, D, dtype=torch.int64, device=device)
        self.row_slot = {t: torch.zeros(spec.table_sizes[t], **i32) for t in self.owned}
        self.touched = torch.zeros(slots, 2, **i32)
        self.touched_count = torch.zeros(1, **i32)
        if self.table_opt == 1:
            self.table_state = {t: torch.zeros(spec.table_sizes[t], **f32) for t in self.owned}
        elif self.table_opt == 2:
            self.table_state = {t: torch.zeros(2, spec.table_sizes[t], D, **f32) for t in self.owned}
        else:
            self.table_state = {}
        # ---- dense parameters: flat fp32 [W, b] per layer, bottom then top --------------------
        bdims = [spec.num_dense] + list(spec.bottom)
        tdims = [self.inter_w] + list(spec.top)
        self.layers = [("bottom", i, bdims[i], bdims[i + 1]) for i in range(len(spec.bottom))] + \
                      [("top", i, tdims[i], tdims[i + 1]) for i in range(len(spec.top))]
        self.nb = len(spec.bottom)
        off, self.slices = 0, []
        for _, _, k, n in self.layers:
            self.slices.append((off, off + n * k, off + n * k + n))
            off += n * k + n
        self.n_params = (off + 3) // 4 * 4
        self.params = torch.zeros(self.n_params, **f32)
        self.grads = torch.zeros(self.n_params, **f32)
        self.momentum_buf: Optional[torch.Tensor] = None
        # bf16 operand copies (K padded to 8 for TMA pitch; last layer padded to 16 outputs)
        self.w_bf, self.w_t, self.npad = [], [], []
        for li, (_, _, k, n) in enumerate(self.layers):
            npad = n if n % 16 == 0 else _r8(max(n, 16) + 8) // 16 * 16
            kp = _r8(k)
            self.npad.append(npad)
            self.w_bf.append(torch.zeros(npad, kp, **bf))
            self.w_t.append(torch.zeros(kp, npad, **bf))
        self.bias_pad = [torch.zeros(np_, **f32) for np_ in self.npad]
        # ---- activations ---------------------------------------------------------------------
        B = self.B
        self.feats = torch.zeros(B, self.S, D, **bf)
        self.dfeats = torch.zeros(B, self.S, D, **bf)
        self.idx = torch.zeros(self.C, self.ldb, dtype=torch.int32, device=device)      # transposed
        self.batch_dev = torch.full((1,), B, dtype=torch.int32, device=device)         # read by the owners
        self.act = [torch.zeros(B, self.npad[i], **bf) for i in range(len(self.layers))]
        self.dact = [torch.zeros(B, self.npad[i], **bf) for i in range(len(self.layers))]
        self.zint = torch.zeros(B, self.inter_w, **bf)
        self.dzint = torch.zeros(B, self.inter_w, **bf)
        self.stats = [torch.zeros(2, max(self.npad[i], 64), **f32) for i in range(len(self.layers))]
        self.stats_bottom_last = torch.zeros(2, max(D, 8), **f32)
        self.accum = torch.zeros(8, **f32)
        # deterministic wgrad: per-layer scratch for the batch splits (at most 8 per layer, 192 MB in total);
        # the splits of all layers are added in order by one launch at the end of the backward pass
        self.wgrad_ws: List[torch.Tensor] = []
        self._wgrad_splits = [1] * len(self.layers)
        if self.deterministic:
            mult = 8
            while mult > 1 and sum(n * k for _, _, k, n in self.layers) * mult > (48 << 20):
                mult //= 2
            pool = torch.empty(sum((n * k * mult + 3) // 4 * 4 for _, _, k, n in self.layers), **f32)
            off = 0
            for _, _, k, n in self.layers:
                self.wgrad_ws.append(pool[off:off + n * k * mult])
                off += (n * k * mult + 3) // 4 * 4
        groups = sum((n + 127) // 128 for _, _, _, n in self.layers) + 1
        self.colsum_ws = torch.empty(32 * groups * 128, **f32) if self.deterministic else None
        self.colsum_counters = torch.zeros(groups, dtype=torch.int32, device=device) if self.deterministic else None
        self.ones32 = torch.ones(D, **f32)
        self.zeros32 = torch.zeros(D, **f32)
        # peers (filled by the owner of the world: symmetric memory or LocalWorld)
        self.peer_idx: List[int] = [self.idx.data_ptr()]
        self.peer_feats: List[int] = [self.feats.data_ptr()]
        self.peer_dfeats: List[int] = [self.dfeats.data_ptr()]
        self.peer_batch: List[int] = [B]
        self.peer_batch_ptr: List[int] = [self.batch_dev.data_ptr()]
        self.step_dev = torch.zeros(1, **f32)          # optimizer step count for the tables (Adam bias correction)

    # ------------------------------------------------------------------ parameters
    def weight(self, li: int, grad: bool = False) -> torch.Tensor:
        _, _, k, n = self.layers[li]
        a, b, _ = self.slices[li]
        return (self.grads if grad else self.params)[a:b].view(n, k)

    def bias(self, li: int, grad: bool = False) -> torch.Tensor:
        _, b, c = self.slices[li]
        return (self.grads if grad else self.params)[b:c]

    def load_from_module(self, module, plan=None) -> None:
        """``plan`` (:class:`raydp_b200.torch.lowering.DLRMPlan`) names the layers of a user-defined module;
        without it the module is ``raydp_b200.models.DLRM`` (``bottom`` / ``top`` / ``embeddings`` lists)."""
        if plan is not None:
            bottom, top, embs = plan.modules(module)
        else:
            bottom, top, embs = list(module.bottom), list(module.top), list(module.embeddings)
        with torch.no_grad():
            for li, lin in enumerate(list(bottom) + list(top)):
                w = self.weight(li)
                if w.shape[1] != lin.in_features:          # first top layer: interaction width padded to 16
                    w.zero_()
                w[:, :lin.in_features].copy_(lin.weight)
                self.bias(li).copy_(lin.bias)
            for t in self.owned:
                self.tables[t].copy_(embs[t].weight)

    def dense_state_dict(self) -> Dict[str, torch.Tensor]:
        out = {}
        for li, (part, i, _, _) in enumerate(self.layers):
            k = self.inter_w_model if li == self.nb else self.layers[li][2]
            out[f"{part}.{i}.weight"] = self.weight(li)[:, :k].detach().cpu().clone()
            out[f"{part}.{i}.bias"] = self.bias(li).detach().cpu().clone()
        return out

    def table_state_dict(self) -> Dict[str, torch.Tensor]:
        return {f"embeddings.{t}.weight": w.detach().cpu().clone() for t, w in self.tables.items()}

    # ------------------------------------------------------------------ pieces of a step
    def _refresh_weights(self) -> None:
        """bf16 GEMM operands (plain and transposed) and padded biases of every layer, one launch."""
        L = range(len(self.layers))
        ops.cast_weight_many([self.weight(li) for li in L], self.w_bf, self.w_t, [self.bias(li) for li in L],
                             self.bias_pad)

    def _wgrad(self, g: torch.Tensor, inp: torch.Tensor, li: int) -> None:
        _, _, k, n = self.layers[li]
        if self.deterministic:
            self._wgrad_splits[li] = ops.gemm_wgrad_partial(g, inp, self.weight(li, True), self.wgrad_ws[li],
                                                            mw=n, nw=k)
        else:
            ops.gemm_wgrad(g, inp, self.weight(li, True), mw=n, nw=k)

    def _dgrad(self, g: torch.Tensor, li: int, prev: int, B: int, k: int) -> None:
        """dact[prev] = (g @ W_li) masked by ReLU of act[prev]; bias gradient of `prev` from the tensor-core column
        sums (atomic across CTAs) or, deterministic, left to :meth:`_bias_grads`."""
        if self.deterministic:
            ops.gemm_tn(g, self.w_t[li], self.dact[prev][:B], aux=self.act[prev][:B], aux_relu_mask=True,
                        n=self.npad[prev], k=k)
            return
        self.stats[prev].zero_()
        ops.gemm_tn(g, self.w_t[li], self.dact[prev][:B], aux=self.act[prev][:B], aux_relu_mask=True,
                    stats_mode=1, stats=self.stats[prev], n=self.npad[prev], k=k)
        self.bias(prev, True).copy_(self.stats[prev][0, :self.layers[prev][3]])

    def _bias_grads(self, layers: Sequence[int], B: int, extra=None) -> None:
        """Ordered column sums of the layers' output gradients -> bias gradients (one launch)."""
        mats = [self.dact[li][:B] for li in layers]
        widths = [self.layers[li][3] for li in layers]
        outs = [self.bias(li, True) for li in layers]
        if extra is not None:
            mats.append(extra[0]); widths.append(extra[1]); outs.append(extra[2])
        if mats:
            ops.colsum_ordered(mats, widths, outs, self.colsum_ws, self.colsum_counters)

    def _gemm_fwd(self, li: int, x: torch.Tensor, out: torch.Tensor, relu: bool) -> None:
        ops.gemm_tn(x, self.w_bf[li], out, bias=self.bias_pad[li], relu=relu, n=self.npad[li],
                    k=self.w_bf[li].shape[1])

    def bottom_forward(self, dense: torch.Tensor) -> None:
        """dense [B, kpad] bf16 -> feats[:, 0, :]"""
        B = dense.shape[0]
        x = dense
        for li in range(self.nb):
            last = li == self.nb - 1
            out = self.feats.view(self.B, self.S * self.D)[:B, :self.D] if last else self.act[li][:B]
            self._gemm_fwd(li, x, out, relu=True)
            x = out

    def set_indices(self, idx: torch.Tensor) -> None:
        """idx [B, C] (any integer dtype) -> this rank's transposed int32 index matrix and batch count."""
        B = idx.shape[0]
        if B:
            self.idx[:, :B].copy_(idx.t())
        self.batch_dev.fill_(B)

    def _owned_args(self):
        return ([self.tables[t].data_ptr() for t in self.owned], [t + 1 for t in self.owned],
                [self.col0[t] for t in self.owned], [self.lens[t] for t in self.owned])

    def embedding_forward(self) -> None:
        """Pool and push rows of the tables this rank owns into every rank's feature slab."""
        if not self.owned:
            return
        tables, slots, cols, lens = self._owned_args()
        ops.embedding_bag_fwd(tables, slots, cols, lens, self.peer_idx, self.peer_feats, self.peer_batch,
                              self.peer_batch_ptr, self.ldb, self.S, self.D, self.mean)

    def top_forward_backward(self, target: Optional[torch.Tensor], B: int, train: bool) -> torch.Tensor:
        ops.dot_interaction_fwd(self.feats[:B], self.zint[:B])
        x = self.zint[:B]
        n_layers = len(self.layers)
        for li in range(self.nb, n_layers):
            last = li == n_layers - 1
            self._gemm_fwd(li, x, self.act[li][:B], relu=not last)
            x = self.act[li][:B]
        logits = self.act[-1][:B]
        if not train:
            if target is not None:
                ops.bce_logits(logits, target, None, self.accum, 0.0)
            return logits
        self.grads.zero_()
        dl = self.dact[-1][:B]
        ops.bce_logits(logits, target, dl, self.accum, 1.0 / B,
                       None if self.deterministic else self.bias(n_layers - 1, True))
        # ---- top backward ----
        for li in range(n_layers - 1, self.nb - 1, -1):
            inp = self.zint[:B] if li == self.nb else self.act[li - 1][:B]
            g = self.dact[li][:B]
            self._wgrad(g, inp, li)
            if li > self.nb:
                # dgrad with the ReLU mask of the previous activation fused in the epilogue; the
                # tensor-core column sums of the result are that layer's bias gradient
                self._dgrad(g, li, li - 1, B, self.npad[li])
            else:
                ops.gemm_tn(g, self.w_t[li], self.dzint[:B], n=self.inter_w, k=self.npad[li])
        if self.deterministic:
            self._bias_grads(range(self.nb, n_layers), B)
        ops.dot_interaction_bwd(self.feats[:B], self.dzint[:B], self.dfeats[:B])
        return logits

    def embedding_backward(self, lr: float, grad_scale: float, lr_dev: Optional[torch.Tensor] = None,
                           step_dev: Optional[torch.Tensor] = None) -> None:
        if not self.owned:
            return
        tables, slots, cols, lens = self._owned_args()
        a = self.table_opt_args
        state = [self.table_state[t].data_ptr() for t in self.owned] if self.table_state else None
        strides = [self.spec.table_sizes[t] * self.D for t in self.owned] if self.table_opt == 2 else None
        ops.embedding_bag_bwd(tables, slots, cols, lens, self.peer_idx, self.peer_dfeats, self.peer_batch,
                              self.peer_batch_ptr, self.ldb, self.S, self.D, self.mean, grad_scale,
                              self.acc, [self.row_slot[t].data_ptr() for t in self.owned],
                              self.touched, self.touched_count,
                              state, strides, self.table_opt, lr, a["eps"], a["beta1"], a["beta2"],
                              a["weight_decay"], lr_dev, step_dev)

    def bottom_backward(self, dense: torch.Tensor) -> None:
        B = dense.shape[0]
        last = self.nb - 1
        pitch = self.S * self.D
        dy = self.dfeats.view(self.B, pitch)[:B, :self.D]
        z = self.feats.view(self.B, pitch)[:B, :self.D]
        # ReLU backward of the bottom output (in place on the slab view) + its bias gradient
        ops.bn_relu_bwd(dy, z, dy, self.stats_bottom_last, self.zeros32, self.ones32, self.ones32,
                        self.zeros32, self.zeros32, self.bias(last, True), has_bn=False)
        g = dy
        for li in range(last, -1, -1):
            inp = dense if li == 0 else self.act[li - 1][:B]
            self._wgrad(g, inp, li)
            if li > 0:
                prev = li - 1
                self._dgrad(g, li, prev, B, self.npad[li] if li != last else self.D)
                g = self.dact[prev][:B]
        if self.deterministic:
            self._bias_grads(range(0, last), B, extra=(dy, self.D, self.bias(last, True)))
            L = range(len(self.layers))          # top and bottom layers: one ordered reduction for all of them
            ops.wgrad_reduce_many(self.wgrad_ws, self._wgrad_splits, [self.weight(li, True) for li in L])

    def sgd(self, lr: float, scale: float = 1.0, momentum: float = 0.0, weight_decay: float = 0.0,
            lr_dev: Optional[torch.Tensor] = None, step_dev: Optional[torch.Tensor] = None) -> None:
        if momentum and self.momentum_buf is None:
            self.momentum_buf = torch.zeros_like(self.params)
        ops.sgd_step(self.params, self.grads, self.momentum_buf if momentum else None, lr, momentum, 0.0,
                     weight_decay, False, False, scale, lr_dev, step_dev)


class LocalWorld:
    """W model-parallel ranks on one device, stepped in lock step (single-GPU tests / emulation)."""

    def __init__(self, spec: DLRMSpec, batch_per_rank: int, world: int, device: torch.device,
                 table_optimizer: str = "sgd", table_opt_args: Optional[dict] = None,
                 deterministic: Optional[bool] = None):
        self.world = world
        self.ranks = [FusedDLRM(spec, batch_per_rank, device, r, world, table_optimizer, table_opt_args,
                                deterministic) for r in range(world)]
        for r in self.ranks:
            r.peer_idx = [q.idx.data_ptr() for q in self.ranks]
            r.peer_feats = [q.feats.data_ptr() for q in self.ranks]
            r.peer_dfeats = [q.dfeats.data_ptr() for q in self.ranks]
            r.peer_batch = [q.B for q in self.ranks]
            r.peer_batch_ptr = [q.batch_dev.data_ptr() for q in self.ranks]

    def load_from_module(self, module, plan=None) -> None:
        for r in self.ranks:
            r.load_from_module(module, plan)

    def train_step(self, dense: Sequence[torch.Tensor], idx: Sequence[torch.Tensor],
                   target: Sequence[torch.Tensor], lr: float) -> None:
        """Ranks may hold different batch sizes (0 included): every owner reads them from ``batch_dev``."""
        W = self.world
        for r, d, i in zip(self.ranks, dense, idx):
            r._refresh_weights()
            r.set_indices(i)
            r.step_dev.add_(1.0)
            if d.shape[0]:
                r.bottom_forward(d)
        for r in self.ranks:                       # (flag round 1 in the multi-GPU engine)
            r.embedding_forward()
        for r, d, t in zip(self.ranks, dense, target):   # (flag round 2)
            if d.shape[0]:
                r.top_forward_backward(t, d.shape[0], True)
            else:
                r.grads.zero_()
        for r in self.ranks:                       # (flag round 3)
            r.embedding_backward(lr, 1.0 / W, None, r.step_dev)
        for r, d in zip(self.ranks, dense):
            if d.shape[0]:
                r.bottom_backward(d)
        # dense gradients: average over ranks (all-reduce), identical SGD on every replica
        total = torch.stack([r.grads for r in self.ranks]).sum(0)
        for r in self.ranks:
            r.grads.copy_(total)
            r.sgd(lr, 1.0 / W)

    def predict(self, dense: Sequence[torch.Tensor], idx: Sequence[torch.Tensor]) -> List[torch.Tensor]:
        for r, d, i in zip(self.ranks, dense, idx):
            r._refresh_weights()
            r.set_indices(i)
            r.bottom_forward(d)
        for r in self.ranks:
            r.embedding_forward()
        return [r.top_forward_backward(None, d.shape[0], False)[:, 0].float().clone()
                for r, d in zip(self.ranks, dense)]


class DistributedDLRM:
    """One process per GPU: symmetric slabs for the embedding exchange, flag rounds between the
    phases, peer-memory all-reduce fused with SGD for the dense gradients."""

    def __init__(self, spec: DLRMSpec, batch: int, device: torch.device, group=None, momentum: float = 0.0,
                 weight_decay: float = 0.0, use_graph: bool = True, table_optimizer: str = "sgd",
                 table_opt_args: Optional[dict] = None, deterministic: Optional[bool] = None):
        import os
        import torch.distributed as dist
        self.use_graph = use_graph and os.environ.get("RAYDP_B200_DLRM_GRAPH", "1") != "0"
        self._graphs = [None, None]
        self._eager_steps = [0, 0]
        self._graph_launches = 0
        self._in_dense = self._in_idx = self._in_target = None
        self._lr_host = None
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.core = FusedDLRM(spec, batch, device, self.rank, self.world, table_optimizer, table_opt_args,
                              deterministic)
        self.device = device
        self.momentum, self.weight_decay = momentum, weight_decay
        self.step_count = 0
        self.lr_dev = torch.zeros(1, dtype=torch.float32, device=device)
        self.step_dev = torch.zeros(1, dtype=torch.float32, device=device)
        self.sym = None
        self.reducer = None
        if self.world > 1:
            from raydp_b200.comm import P2PReducer, SymmetricMemory, tensor_from_ptr
            c = self.core
            B, S, D = c.B, c.S, c.D
            idx_bytes = (c.C * c.ldb * 4 + 255) // 256 * 256
            slab_bytes = (B * S * D * 2 + 255) // 256 * 256
            count_off = idx_bytes + 2 * slab_bytes             # this rank's batch of the step (int32)
            per_parity = count_off + 256
            self.flag_off = 2 * per_parity
            self.sym = SymmetricMemory(self.flag_off + 256, group)
            self.barrier_epoch = torch.zeros(1, dtype=torch.int32, device=device)
            self.views, self.peer_ptrs = [], []
            for p in range(2):
                base = p * per_parity
                local = self.sym.local + base
                self.views.append((tensor_from_ptr(local, (c.C, c.ldb), torch.int32, device),
                                   tensor_from_ptr(local + idx_bytes, (B, S, D), torch.bfloat16, device),
                                   tensor_from_ptr(local + idx_bytes + slab_bytes, (B, S, D), torch.bfloat16, device),
                                   tensor_from_ptr(local + count_off, (1,), torch.int32, device)))
                self.peer_ptrs.append(([b + base for b in self.sym.bases],
                                       [b + base + idx_bytes for b in self.sym.bases],
                                       [b + base + idx_bytes + slab_bytes for b in self.sym.bases],
                                       [b + base + count_off for b in self.sym.bases]))
            self.core.peer_batch = [B] * self.world
            self.reducer = P2PReducer(self.core.grads, device, group, algo="auto")
            dist.broadcast(self.core.params, src=0)

    def _select(self, parity: int) -> None:
        c = self.core
        c.idx, c.feats, c.dfeats, c.batch_dev = self.views[parity]
        c.peer_idx, c.peer_feats, c.peer_dfeats, c.peer_batch_ptr = self.peer_ptrs[parity]

    def _barrier(self) -> None:
        ops.comm_barrier(self.sym.bases, self.flag_off, self.rank, self.world, self.barrier_epoch)

    def _step_body(self, dense: torch.Tensor, idx: torch.Tensor, target: torch.Tensor, parity: int) -> None:
        """One training step as a pure stream of kernels: no host reads, no allocation, the learning
        rate and the step count come from device scalars -- so it can be captured in a CUDA graph."""
        c = self.core
        B = dense.shape[0]
        multi = self.world > 1
        if multi:
            self._select(parity)
        ops.add_scalar(self.step_dev, 1.0)
        c._refresh_weights()
        c.set_indices(idx)
        if B:
            c.bottom_forward(dense)
        if multi:
            self._barrier()
        c.embedding_forward()
        if multi:
            self._barrier()
        if B:
            c.top_forward_backward(target, B, True)
        else:                   # this rank ran out of data: it still serves its tables and joins the reduction
            c.grads.zero_()
        if multi:
            self._barrier()
        c.embedding_backward(0.0, 1.0 / self.world, self.lr_dev, self.step_dev)
        if B:
            c.bottom_backward(dense)
        if multi:
            from raydp_b200.comm import make_opt_args
            opt = make_opt_args(2, c.params, c.momentum_buf if self.momentum else None, None,
                                weight_decay=self.weight_decay, momentum=self.momentum, lr_dev=self.lr_dev,
                                step_dev=self.step_dev)
            self.reducer.allreduce_step_(opt, 1.0 / self.world)
        else:
            c.sgd(0.0, 1.0, self.momentum, self.weight_decay, self.lr_dev, self.step_dev)

    def train_step(self, dense: torch.Tensor, idx: torch.Tensor, target: torch.Tensor, lr: float) -> None:
        """Full batches replay a captured CUDA graph (one per slab parity; ~100 launches of a
        latency-bound step become one), after one eager step per parity has warmed lazy state.
        Inputs are staged into fixed device buffers so the graphs can read them."""
        c = self.core
        B = dense.shape[0]
        assert B <= c.B, f"batch {B} exceeds the capacity {c.B} the workspaces were sized for"
        parity = self.step_count & 1
        self.step_count += 1
        if self.momentum and c.momentum_buf is None:
            c.momentum_buf = torch.zeros_like(c.params)
        if lr != self._lr_host:
            self.lr_dev.fill_(lr)
            self._lr_host = lr
        if not (self.use_graph and B == c.B):
            self._step_body(dense, idx, target, parity)
            return
        if self._in_dense is None:
            self._in_dense = torch.zeros_like(dense)
            self._in_idx = torch.zeros_like(idx)
            self._in_target = torch.zeros_like(target)
        self._in_dense.copy_(dense)
        self._in_idx.copy_(idx)
        self._in_target.copy_(target)
        graph = self._graphs[parity]
        if graph is not None:
            graph.replay()
            ops.add_launches(self._graph_launches)
        elif self._eager_steps[parity] < 1:
            self._step_body(self._in_dense, self._in_idx, self._in_target, parity)
            self._eager_steps[parity] += 1
        else:
            graph = torch.cuda.CUDAGraph()
            before = ops.launch_count()
            with torch.cuda.graph(graph):
                self._step_body(self._in_dense, self._in_idx, self._in_target, parity)
            self._graph_launches = ops.launch_count() - before
            self._graphs[parity] = graph
            graph.replay()

    def eval_step(self, dense: torch.Tensor, idx: torch.Tensor, target: Optional[torch.Tensor]) -> torch.Tensor:
        c = self.core
        B = dense.shape[0]
        multi = self.world > 1
        if multi:
            self._select(self.step_count & 1)
        self.step_count += 1
        c._refresh_weights()
        c.set_indices(idx)
        if B:
            c.bottom_forward(dense)
        if multi:
            self._barrier()
        c.embedding_forward()
        if multi:
            self._barrier()
        out = c.top_forward_backward(target, B, False) if B else c.act[-1][:0]
        if multi:
            self._barrier()          # keeps the parity protocol's three rounds per step
        return out

    def full_state_dict(self) -> Dict[str, torch.Tensor]:
        """Rank 0 gets the whole model (dense + every table shard); other ranks get {}."""
        import torch.distributed as dist
        state = self.core.dense_state_dict() if self.rank == 0 else {}
        tables = self.core.table_state_dict()
        if self.world > 1:
            gathered = [None] * self.world if self.rank == 0 else None
            dist.gather_object(tables, gathered, dst=0)
            if self.rank == 0:
                for part in gathered:
                    state.update(part)
        else:
            state.update(tables)
        return state

    def close(self) -> None:
        self._graphs = [None, None]
        if self.reducer is not None:
            self.reducer.close()
        if self.sym is not None:
            self.sym.close()
