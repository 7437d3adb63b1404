"""GPU step engine of ``TorchEstimator``: same observable results as the reference loop
(reference ``torch/estimator.py:272-310``), hot path on hand-written sm_100a kernels.

Per training step (one CUDA-graph replay once shapes are warm):
ingest-pack (K5) -> fused forward/backward (K6, K7, K8) -> data-parallel gradient reduction fused
with the optimizer (K1/K2 over NVLink peer memory when world_size > 1) -> asynchronous 32-byte
device->host copy of the running loss/metric sums.  There is no ``loss.item()``: the sums stay on
the device and are read once per epoch, giving the same "mean of per-batch means" the reference
reports.
"""
from __future__ import annotations

import time
from typing import Any, Dict, Optional, Tuple

import torch
import torch.distributed as dist

from raydp_b200 import ops
from raydp_b200.ops.fused_mlp import FusedMLP
from raydp_b200.ops.ingest import ArrowGpuLoader
from raydp_b200.utils.profiling import nvtx_range

_LOSSES = {torch.nn.MSELoss: "mse", torch.nn.SmoothL1Loss: "smooth_l1", torch.nn.BCEWithLogitsLoss: "bce_logits"}


def _loss_kind(criterion) -> Optional[str]:
    for cls, name in _LOSSES.items():
        if type(criterion) is cls:
            if getattr(criterion, "reduction", "mean") != "mean":
                return None
            if name == "smooth_l1" and float(getattr(criterion, "beta", 1.0)) != 1.0:
                return None
            if name == "bce_logits" and (criterion.pos_weight is not None or criterion.weight is not None):
                return None
            return name
    return None


def _optimizer_kind(optimizer) -> Optional[str]:
    if len(optimizer.param_groups) != 1:
        return None
    g = optimizer.param_groups[0]
    if type(optimizer) in (torch.optim.Adam, torch.optim.AdamW):
        if g.get("amsgrad") or g.get("maximize"):
            return None
        return "adamw" if type(optimizer) is torch.optim.AdamW else "adam"
    if type(optimizer) is torch.optim.SGD:
        if g.get("maximize"):
            return None
        return "sgd"
    return None


def select_fused(config, model, optimizer, criterion) -> Tuple[Optional[Any], str]:
    """Decide whether one of the fully fused engines can run this (model, loss, optimizer).

    Returns ``(plan, "")`` -- an ``MLPPlan`` for :class:`FusedEngine` or a ``DLRMPlan`` for
    :class:`DLRMEngine` -- or ``(None, why)``; the caller then uses the generic GPU engine.  The model
    may be ANY ``nn.Module`` (the reference takes arbitrary modules, ``torch/estimator.py:180-187``):
    the plan is derived from the module itself by :mod:`raydp_b200.torch.lowering`."""
    from raydp_b200.torch.lowering import lower_module
    plan, why = lower_module(model)
    if plan is None:
        return None, why
    if not ops.kernels_available():
        raise RuntimeError("CUDA worker without librdp_kernels.so: run __graft_entry__.build() first")
    if plan.kind == "dlrm":
        if type(criterion) is not torch.nn.BCEWithLogitsLoss or _loss_kind(criterion) != "bce_logits":
            return None, "the fused DLRM path needs BCEWithLogitsLoss(mean)"
        if _optimizer_kind(optimizer) != "sgd" or optimizer.param_groups[0].get("nesterov"):
            return None, ("the fused DLRM path needs torch.optim.SGD for the dense layers (the tables choose "
                          "SGD / row-wise Adagrad / Adam through embedding_optimizer=)")
        n_cols = len(config.get("feature_columns") or [])
        if n_cols != plan.spec.num_dense + plan.spec.index_columns:
            return None, ("feature_columns must be the dense columns followed by one column per embedding table "
                          "(bag_sizes[t] columns for a multi-hot table)")
        return plan, ""
    if isinstance(config.get("feature_types"), (list, tuple)):
        return None, "per-column feature dtypes feed model(*columns): generic path"
    if _loss_kind(criterion) is None:
        return None, f"loss {type(criterion).__name__} is not one of MSE / SmoothL1(beta=1) / BCEWithLogits (mean)"
    if _optimizer_kind(optimizer) is None:
        return None, f"optimizer {type(optimizer).__name__} is not Adam / AdamW / SGD with one param group"
    return plan, ""


def fused_supported(config, model, optimizer, criterion) -> Tuple[bool, str]:
    plan, why = select_fused(config, model, optimizer, criterion)
    return plan is not None, why


def _seed_moments(optimizer, named_params: Dict[str, torch.nn.Parameter], targets, m, v, step_dev) -> None:
    """Carry optimizer state that came in with a pre-trained optimizer instance (Adam moments, SGD
    momentum buffers, step count) over into the flat fused buffers.  ``targets``: iterable of
    ``(parameter name, offset, numel)`` into ``m`` / ``v``."""
    step = 0.0
    with torch.no_grad():
        for name, off, n in targets:
            p = named_params.get(name)
            st = optimizer.state.get(p) if p is not None else None
            if not st:
                continue
            if "exp_avg" in st:
                m[off:off + n].copy_(st["exp_avg"].reshape(-1))
                v[off:off + n].copy_(st["exp_avg_sq"].reshape(-1))
            elif st.get("momentum_buffer") is not None:
                m[off:off + n].copy_(st["momentum_buffer"].reshape(-1))
                step = max(step, 1.0)
            if "step" in st:
                step = max(step, float(st["step"]))
        if step:
            step_dev.fill_(step)


class FusedEngine:
    name = "fused"

    def __init__(self, config, sess, model, optimizer, criterion, scheduler, metrics, plan=None):
        self.config = config
        self.sess = sess
        self.device = sess.device
        self.optimizer = optimizer
        self.scheduler = scheduler
        self.metrics = metrics
        self.world = sess.world_size
        self.batch = int(config["batch_size"])
        self.opt_kind = _optimizer_kind(optimizer)
        if plan is None:
            plan, why = select_fused(config, model, optimizer, criterion)
            if plan is None:
                raise RuntimeError(f"FusedEngine cannot run this model: {why}")
        self.plan = plan
        spec = plan.spec
        self.fused = FusedMLP(spec, self.batch, self.device, loss=_loss_kind(criterion))
        self.fused.load_state_dict({k: v.to(self.device) for k, v in model.state_dict().items()})
        if self.world > 1 and dist.is_initialized():
            dist.broadcast(self.fused.params, src=0)       # identical replicas
            for t in self.fused.running_mean + self.fused.running_var:
                dist.broadcast(t, src=0)
        n = self.fused.n_params
        f32 = dict(dtype=torch.float32, device=self.device)
        self.m = torch.zeros(n, **f32)
        self.v = torch.zeros(n, **f32)
        self.step_dev = torch.zeros(1, **f32)
        self.lr_dev = torch.zeros(1, **f32)
        if optimizer.state:
            sl = self.fused._slices
            L = self.fused.L
            names = []
            for l in range(L):
                names += [(f"{spec.linear_names[l]}.weight", f"W{l}"), (f"{spec.linear_names[l]}.bias", f"b{l}")]
                if spec.batch_norm:
                    names += [(f"{spec.bn_names[l]}.weight", f"g{l}"), (f"{spec.bn_names[l]}.bias", f"be{l}")]
            names += [(f"{spec.linear_names[L]}.weight", "Wh"), (f"{spec.linear_names[L]}.bias", "bh")]
            _seed_moments(optimizer, dict(model.named_parameters()),
                          [(pn, sl[k][0], int(torch.tensor(sl[k][1]).prod())) for pn, k in names],
                          self.m, self.v, self.step_dev)
        self._lr_host = None
        self._graphs: Dict[Tuple[int, int], torch.cuda.CUDAGraph] = {}
        self._graph_launches = 0
        self._eager_steps = 0
        self.use_graph = bool(config.get("extra", {}).get("cuda_graph", True))
        self._accum_host = torch.zeros(8, dtype=torch.float32).pin_memory()
        self.reducer = None
        self.group_reducers = None
        if self.world > 1:
            from raydp_b200.parallel import make_gradient_reducer
            extra = config.get("extra", {})
            self.reducer = make_gradient_reducer(self.fused.grads, self.device, extra)
            ranges = self.fused.group_ranges()
            if self.reducer.kind == "p2p" and extra.get("overlap_allreduce", True) and len(ranges) > 1 \
                    and all(lo % 4 == 0 and hi % 4 == 0 for lo, hi in ranges):
                # one reducer per layer: backward finishes the layers last-first, and each layer's
                # all-reduce + optimizer runs on a side stream while the earlier layers are still in backward
                from raydp_b200.comm import P2PReducer
                self.reducer.close()
                self.reducer = None
                self.group_ranges = ranges
                self.group_reducers = [P2PReducer(self.fused.grads[lo:hi], self.device) for lo, hi in ranges]
                self.comm_stream = torch.cuda.Stream(device=self.device, priority=-1)
        self.gpu_launches = 0
        self.d2h_bytes = 0

    # ------------------------------------------------------------------ one step
    def _sync_lr(self) -> None:
        lr = float(self.optimizer.param_groups[0]["lr"])
        if lr != self._lr_host:
            self.lr_dev.fill_(lr)
            self._lr_host = lr

    def _update(self) -> None:
        g = self.optimizer.param_groups[0]
        ops.add_scalar(self.step_dev, 1.0)
        scale = 1.0 / self.world
        adam = self.opt_kind in ("adam", "adamw")
        if self.reducer is not None and self.reducer.kind == "p2p":
            # K1/K2: pull the peers' gradients over NVLink, scale, and apply the optimizer in ONE kernel
            from raydp_b200.comm import make_opt_args
            if adam:
                opt = make_opt_args(1, self.fused.params, self.m, self.v, betas=g["betas"], eps=g["eps"],
                                    weight_decay=g["weight_decay"], adamw=self.opt_kind == "adamw",
                                    lr_dev=self.lr_dev, step_dev=self.step_dev)
            else:
                opt = make_opt_args(2, self.fused.params, self.m if g["momentum"] else None, None,
                                    weight_decay=g["weight_decay"], momentum=g["momentum"],
                                    dampening=g["dampening"], nesterov=g["nesterov"],
                                    lr_dev=self.lr_dev, step_dev=self.step_dev)
            self.reducer.allreduce_step_(opt, scale)
            return
        if self.reducer is not None:
            self.reducer.allreduce_()                      # NCCL baseline: sum over ranks, in place
        if adam:
            b1, b2 = g["betas"]
            ops.adam_step(self.fused.params, self.fused.grads, self.m, self.v, 0.0, b1, b2, g["eps"],
                          g["weight_decay"], 1, scale, self.opt_kind == "adamw", self.lr_dev, self.step_dev)
        else:
            ops.sgd_step(self.fused.params, self.fused.grads, self.m if g["momentum"] else None, 0.0,
                         g["momentum"], g["dampening"], g["weight_decay"], g["nesterov"], False, scale,
                         self.lr_dev, self.step_dev)

    def _group_opt(self, lo: int, hi: int):
        from raydp_b200.comm import make_opt_args
        g = self.optimizer.param_groups[0]
        p, m, v = self.fused.params[lo:hi], self.m[lo:hi], self.v[lo:hi]
        if self.opt_kind in ("adam", "adamw"):
            return make_opt_args(1, p, m, v, betas=g["betas"], eps=g["eps"], weight_decay=g["weight_decay"],
                                 adamw=self.opt_kind == "adamw", lr_dev=self.lr_dev, step_dev=self.step_dev)
        return make_opt_args(2, p, m if g["momentum"] else None, None, weight_decay=g["weight_decay"],
                             momentum=g["momentum"], dampening=g["dampening"], nesterov=g["nesterov"],
                             lr_dev=self.lr_dev, step_dev=self.step_dev)

    def _step_body_overlapped(self, x, y) -> None:
        """Data-parallel step with the gradient reduction off the critical path: when the last gradient of a
        layer is final, its all-reduce + optimizer kernel goes to the communication stream (behind an event)
        and backward continues; only the first layer's small reduction is left after backward ends."""
        compute = torch.cuda.current_stream(self.device)
        comm = self.comm_stream
        scale = 1.0 / self.world
        ops.add_scalar(self.step_dev, 1.0)

        def ready(l: int) -> None:
            ev = torch.cuda.Event()
            ev.record(compute)
            comm.wait_event(ev)
            lo, hi = self.group_ranges[l]
            with torch.cuda.stream(comm):
                self.group_reducers[l].allreduce_step_(self._group_opt(lo, hi), scale)
        self.fused.train_step(x, y, on_group_ready=ready)
        done = torch.cuda.Event()
        done.record(comm)
        compute.wait_event(done)
        ops.publish_host(self.fused.accum, self._accum_host)

    def _step_body(self, x, y) -> None:
        if self.group_reducers is not None:
            return self._step_body_overlapped(x, y)
        self.fused.train_step(x, y)
        self._update()
        # the step's result (running loss / metric sums, 32 bytes) goes to pinned host memory from inside
        # the step -- part of the replayed graph -- by a device store, not a memcpy (ops.publish_host)
        ops.publish_host(self.fused.accum, self._accum_host)

    def _train_step(self, x, y) -> None:
        """Full batches replay a captured CUDA graph.  The loader alternates between two output slots
        (batch i+1 is copied and packed while step i runs), so there is one graph per input buffer."""
        self._sync_lr()
        full = x.shape[0] == self.batch
        key = (x.data_ptr(), y.data_ptr())
        graph = self._graphs.get(key) if (self.use_graph and full) else None
        if not (self.use_graph and full):
            self._step_body(x, y)
        elif graph is not None:
            graph.replay()
            ops.add_launches(self._graph_launches)
            if self.fused.spec.batch_norm:
                self.fused.num_batches_tracked += 1        # host-side counter: a replay does not run train_step
        elif self._eager_steps < 1:
            self._step_body(x, y)                          # first step warms lazy CUDA state
            self._eager_steps += 1
        elif len(self._graphs) >= 4:
            self._step_body(x, y)                          # a caller that keeps changing buffers: stay eager
        else:
            graph = torch.cuda.CUDAGraph()
            before = ops.launch_count()
            with torch.cuda.graph(graph):
                self._step_body(x, y)                      # x / y are the loader's fixed buffers
            self._graph_launches = ops.launch_count() - before
            self._graphs[key] = graph
            graph.replay()                                 # (the capture pass counted this step already)
        if self.scheduler is not None:
            self.scheduler.step()

    # ------------------------------------------------------------------ epochs
    def _loader(self, shard) -> ArrowGpuLoader:
        c = self.config
        cache = getattr(self, "_loaders", None)
        if cache is None:
            cache = self._loaders = {}
        key = id(shard)
        if key not in cache:
            loader = ArrowGpuLoader(shard, c["feature_columns"], c["label_column"], self.batch,
                                    c["drop_last"], self.device, ld=self.fused.kpad,
                                    null_fill=float(c.get("extra", {}).get("null_fill", float("nan"))))
            loader._prepare()
            if loader.n_features != self.fused.spec.in_features:
                raise ValueError(f"the feature columns provide {loader.n_features} values per row (vector columns "
                                 f"count with their width) but the model's first layer takes "
                                 f"{self.fused.spec.in_features}")
            cache[key] = loader
        return cache[key]

    def _results(self, acc, steps: int) -> Tuple[Dict[str, float], float]:
        count = max(1.0, acc[3])
        builtin = {"MeanAbsoluteError": acc[1] / count, "MeanSquaredError": acc[2] / count,
                   "Accuracy": acc[4] / count, "BinaryAccuracy": acc[4] / count}
        res = {}
        for name in self.metrics.names:
            if name in builtin:
                metric = self.metrics.get(name)
                value = builtin[name]
                if name == "MeanSquaredError" and getattr(metric, "squared", True) is False:
                    value = value ** 0.5
                res[name] = float(value)
        return res, float(acc[5] / max(1, steps))

    def train_epoch(self, shard):
        loader = self._loader(shard)
        extra_metrics = [n for n in self.metrics.names
                         if n not in ("MeanAbsoluteError", "MeanSquaredError", "Accuracy", "BinaryAccuracy")]
        self.fused.accum.zero_()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        start.record()
        steps = samples = 0
        h2d0 = loader.h2d_bytes
        for x, y in loader:
            with nvtx_range("rdp.fused.step"):
                self._train_step(x, y)
            if extra_metrics:
                for n in extra_metrics:
                    self.metrics.get(n).update(self.fused.pred[:x.shape[0]].unsqueeze(1), y.unsqueeze(1))
            self.d2h_bytes += 32                 # written by the step itself (see _step_body)
            steps += 1
            samples += x.shape[0]
        end.record()
        torch.cuda.current_stream().synchronize()
        if self.world > 1:
            from raydp_b200.comm import check_status
            check_status()          # a peer wait that timed out surfaces here -> FailureConfig restarts the group
        dt = time.perf_counter() - t0
        acc = self._accum_host.tolist()
        res, loss = self._results(acc, steps)
        for n in extra_metrics:
            res[n] = float(self.metrics.get(n).compute())
            self.metrics.get(n).reset()
        timing = {"time_this_iter_s": dt, "device_time_s": start.elapsed_time(end) / 1e3,
                  "samples_this_iter": samples, "steps_this_iter": steps, "engine": self.name,
                  "h2d_bytes_this_iter": loader.h2d_bytes - h2d0, "d2h_bytes_this_iter": 32 * steps}
        return res, loss, timing

    def evaluate_epoch(self, shard):
        loader = self._loader(shard)
        self.fused.accum.zero_()
        steps = 0
        extra_metrics = [n for n in self.metrics.names
                         if n not in ("MeanAbsoluteError", "MeanSquaredError", "Accuracy", "BinaryAccuracy")]
        for x, y in loader:
            pred = self.fused.eval_step(x, y)
            for n in extra_metrics:
                self.metrics.get(n).update(pred.unsqueeze(1), y.unsqueeze(1))
            steps += 1
        acc = self.fused.accum.tolist()
        res, loss = self._results(acc, steps)
        for n in extra_metrics:
            res[n] = float(self.metrics.get(n).compute())
            self.metrics.get(n).reset()
        return res, loss

    def state_dict(self) -> Dict[str, torch.Tensor]:
        torch.cuda.current_stream().synchronize()
        return self.fused.state_dict()

    def close(self) -> None:
        if self.reducer is not None:
            self.reducer.close()
        for r in self.group_reducers or []:
            r.close()


class DLRMEngine:
    """Estimator engine for :class:`raydp_b200.models.DLRM` (hybrid parallel, see
    ``raydp_b200.parallel.hybrid``): dense columns first, then one categorical column per table."""

    name = "fused-dlrm"

    def __init__(self, config, sess, model, optimizer, criterion, scheduler, metrics, plan=None):
        from raydp_b200.parallel.hybrid import DistributedDLRM
        self.config, self.sess, self.device = config, sess, sess.device
        self.optimizer, self.scheduler, self.metrics = optimizer, scheduler, metrics
        self.batch = int(config["batch_size"])
        g = optimizer.param_groups[0]
        if plan is None:
            plan, why = select_fused(config, model, optimizer, criterion)
            if plan is None or plan.kind != "dlrm":
                raise RuntimeError(f"DLRMEngine cannot run this model: {why}")
        self.plan = plan
        self.spec = plan.spec
        extra = config.get("extra", {}) or {}
        # the tables may use their own sparse optimizer: extra={"embedding_optimizer": "adagrad" | "adam" | "sgd",
        # "embedding_optimizer_args": {"eps": ..., "beta1": ..., "beta2": ..., "weight_decay": ...}}
        self.dlrm = DistributedDLRM(self.spec, self.batch, self.device, momentum=g.get("momentum", 0.0),
                                    weight_decay=g.get("weight_decay", 0.0),
                                    table_optimizer=extra.get("embedding_optimizer", "sgd"),
                                    table_opt_args=extra.get("embedding_optimizer_args"))
        # each rank copies the dense layers and ONLY the tables it owns straight from the host module
        self.dlrm.core.load_from_module(model, plan)
        cols = list(config["feature_columns"])
        self.dense_cols = cols[:self.spec.num_dense]
        self.sparse_cols = cols[self.spec.num_dense:]
        assert len(self.sparse_cols) == self.spec.index_columns, \
            "one categorical column per embedding table (bag_sizes[t] columns for a multi-hot table)"
        self._accum_host = torch.zeros(8, dtype=torch.float32).pin_memory()
        self._loaders = {}

    def _loader(self, shard):
        if id(shard) not in self._loaders:
            c = self.config
            self._loaders[id(shard)] = ArrowGpuLoader(
                shard, self.dense_cols, c["label_column"], self.batch, c["drop_last"],
                self.device, ld=self.dlrm.core.kpad, sparse_columns=self.sparse_cols,
                table_rows=self._column_rows(),
                null_fill=float(c.get("extra", {}).get("null_fill", float("nan"))))
        return self._loaders[id(shard)]

    def _column_rows(self):
        lens = self.spec.bag_sizes or [1] * len(self.spec.table_sizes)
        # bag columns keep negative / NULL entries as padding (-1); single columns wrap into the table
        return [rows if n == 1 else -rows for rows, n in zip(self.spec.table_sizes, lens) for _ in range(n)]

    def _common_steps(self, loader) -> int:
        """Ranks own tables of every other rank's lookups, so all of them take the same number of steps: the
        longest shard decides, ranks that run out of rows join with an empty batch."""
        n = len(loader)
        if self.sess.world_size > 1:
            import torch.distributed as dist
            t = torch.tensor([n], dtype=torch.int64, device=self.device)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            n = int(t.item())
        return n

    def _empty_batch(self):
        core = self.dlrm.core
        return (torch.empty(0, core.kpad, dtype=torch.bfloat16, device=self.device),
                torch.empty(0, core.C, dtype=torch.int32, device=self.device),
                torch.empty(0, dtype=torch.float32, device=self.device))

    def _results(self, acc, steps):
        count = max(1.0, acc[3])
        res = {}
        for name in self.metrics.names:
            if name in ("Accuracy", "BinaryAccuracy"):
                res[name] = float(acc[4] / count)
        return res, float(acc[5] / max(1, steps))

    def train_epoch(self, shard):
        loader = self._loader(shard)
        core = self.dlrm.core
        core.accum.zero_()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        start.record()
        steps = samples = 0
        h2d0 = loader.h2d_bytes
        total_steps = self._common_steps(loader)
        for dense, idx, y in loader:
            with nvtx_range("rdp.dlrm.step"):
                self.dlrm.train_step(dense, idx, y, float(self.optimizer.param_groups[0]["lr"]))
            if self.scheduler is not None:
                self.scheduler.step()
            ops.publish_host(core.accum, self._accum_host)
            steps += 1
            samples += dense.shape[0]
        own_steps = steps
        while steps < total_steps:
            self.dlrm.train_step(*self._empty_batch(), float(self.optimizer.param_groups[0]["lr"]))
            if self.scheduler is not None:
                self.scheduler.step()
            steps += 1
        end.record()
        torch.cuda.current_stream().synchronize()
        if self.sess.world_size > 1:
            from raydp_b200.comm import check_status
            check_status()
        res, loss = self._results(self._accum_host.tolist(), own_steps)
        timing = {"time_this_iter_s": time.perf_counter() - t0, "device_time_s": start.elapsed_time(end) / 1e3,
                  "samples_this_iter": samples, "steps_this_iter": steps, "engine": self.name,
                  "h2d_bytes_this_iter": loader.h2d_bytes - h2d0, "d2h_bytes_this_iter": 32 * steps}
        return res, loss, timing

    def evaluate_epoch(self, shard):
        loader = self._loader(shard)
        core = self.dlrm.core
        core.accum.zero_()
        steps = 0
        total_steps = self._common_steps(loader)
        for dense, idx, y in loader:
            self.dlrm.eval_step(dense, idx, y)
            steps += 1
        own_steps = steps
        while steps < total_steps:
            self.dlrm.eval_step(*self._empty_batch()[:2], None)
            steps += 1
        return self._results(core.accum.tolist(), own_steps)

    def state_dict(self):
        torch.cuda.current_stream().synchronize()
        state = self.dlrm.full_state_dict()
        return self.plan.rename_state(state) if state else state

    def close(self):
        self.dlrm.close()
