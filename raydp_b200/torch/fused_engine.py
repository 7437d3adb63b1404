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


def _is_dlrm(model) -> bool:
    from raydp_b200.models.dlrm import DLRMSpec
    return hasattr(model, "fused_spec") and isinstance(model.fused_spec(), DLRMSpec)


def fused_supported(config, model, optimizer, criterion) -> Tuple[bool, str]:
    if not hasattr(model, "fused_spec"):
        return False, "model does not describe a fusable layer chain (no fused_spec())"
    if _is_dlrm(model):
        if type(criterion) is not torch.nn.BCEWithLogitsLoss or _loss_kind(criterion) != "bce_logits":
            return False, "the fused DLRM path needs BCEWithLogitsLoss(mean)"
        if _optimizer_kind(optimizer) != "sgd" or optimizer.param_groups[0].get("nesterov"):
            return False, "the fused DLRM path needs torch.optim.SGD (sparse embedding updates are SGD)"
        if not ops.kernels_available():
            raise RuntimeError("CUDA worker without librdp_kernels.so: run __graft_entry__.build() first")
        return True, ""
    if isinstance(config.get("feature_types"), (list, tuple)):
        return False, "per-column feature dtypes need the generic path"
    if _loss_kind(criterion) is None:
        return False, f"loss {type(criterion).__name__} is not one of MSE / SmoothL1(beta=1) / BCEWithLogits (mean)"
    if _optimizer_kind(optimizer) is None:
        return False, f"optimizer {type(optimizer).__name__} is not Adam / AdamW / SGD with one param group"
    try:
        model.fused_spec()
    except ValueError as exc:
        return False, str(exc)
    if not ops.kernels_available():
        raise RuntimeError("CUDA worker without librdp_kernels.so: run __graft_entry__.build() first")
    return True, ""


class FusedEngine:
    name = "fused"

    def __init__(self, config, sess, model, optimizer, criterion, scheduler, metrics):
        self.config = config
        self.sess = sess
        self.device = sess.device
        self.optimizer = optimizer
        self.scheduler = scheduler
        self.metrics = metrics
        self.world = sess.world_size
        self.batch = int(config["batch_size"])
        self.opt_kind = _optimizer_kind(optimizer)
        spec = model.fused_spec()
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
        self._lr_host = None
        self._graph: Optional[torch.cuda.CUDAGraph] = None
        self._eager_steps = 0
        self.use_graph = bool(config.get("extra", {}).get("cuda_graph", True))
        self._accum_host = torch.zeros(8, dtype=torch.float32).pin_memory()
        self.reducer = None
        if self.world > 1:
            from raydp_b200.parallel import make_gradient_reducer
            self.reducer = make_gradient_reducer(self.fused.grads, self.device, config.get("extra", {}))
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

    def _step_body(self, x, y) -> None:
        self.fused.train_step(x, y)
        self._update()

    def _train_step(self, x, y) -> None:
        self._sync_lr()
        full = x.shape[0] == self.batch
        if not (self.use_graph and full):
            self._step_body(x, y)
        elif self._graph is not None:
            self._graph.replay()
            ops.add_launches(self._graph_launches)
        elif self._eager_steps < 1:
            self._step_body(x, y)                          # first step warms lazy CUDA state
            self._eager_steps += 1
        else:
            graph = torch.cuda.CUDAGraph()
            before = ops.launch_count()
            with torch.cuda.graph(graph):
                self._step_body(x, y)                      # x / y are the loader's fixed buffers
            self._graph_launches = ops.launch_count() - before
            self._graph = graph
            graph.replay()
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
            cache[key] = ArrowGpuLoader(shard, c["feature_columns"], c["label_column"], self.batch,
                                        c["drop_last"], self.device, ld=self.fused.kpad)
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
            self._train_step(x, y)
            if extra_metrics:
                for n in extra_metrics:
                    self.metrics.get(n).update(self.fused.pred[:x.shape[0]].unsqueeze(1), y.unsqueeze(1))
            # the step's result (running loss / metric sums) goes to pinned host memory, async
            self._accum_host.copy_(self.fused.accum, non_blocking=True)
            self.d2h_bytes += 32
            steps += 1
            samples += x.shape[0]
        end.record()
        torch.cuda.current_stream().synchronize()
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


class DLRMEngine:
    """Estimator engine for :class:`raydp_b200.models.DLRM` (hybrid parallel, see
    ``raydp_b200.parallel.hybrid``): dense columns first, then one categorical column per table."""

    name = "fused-dlrm"

    def __init__(self, config, sess, model, optimizer, criterion, scheduler, metrics):
        from raydp_b200.parallel.hybrid import DistributedDLRM
        self.config, self.sess, self.device = config, sess, sess.device
        self.optimizer, self.scheduler, self.metrics = optimizer, scheduler, metrics
        self.batch = int(config["batch_size"])
        g = optimizer.param_groups[0]
        self.spec = model.fused_spec()
        self.dlrm = DistributedDLRM(self.spec, self.batch, self.device, momentum=g.get("momentum", 0.0),
                                    weight_decay=g.get("weight_decay", 0.0))
        self.dlrm.core.load_from_module(model.to(self.device))
        model.to("cpu")
        cols = list(config["feature_columns"])
        self.dense_cols = cols[:self.spec.num_dense]
        self.sparse_cols = cols[self.spec.num_dense:]
        assert len(self.sparse_cols) == len(self.spec.table_sizes), "one categorical column per embedding table"
        self._accum_host = torch.zeros(8, dtype=torch.float32).pin_memory()
        self._loaders = {}

    def _loader(self, shard):
        if id(shard) not in self._loaders:
            c = self.config
            self._loaders[id(shard)] = ArrowGpuLoader(
                shard, self.dense_cols, c["label_column"], self.batch, c["drop_last"] or self.sess.world_size > 1,
                self.device, ld=self.dlrm.core.kpad, sparse_columns=self.sparse_cols,
                table_rows=list(self.spec.table_sizes))
        return self._loaders[id(shard)]

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
        for dense, idx, y in loader:
            self.dlrm.train_step(dense, idx, y, float(self.optimizer.param_groups[0]["lr"]))
            if self.scheduler is not None:
                self.scheduler.step()
            self._accum_host.copy_(core.accum, non_blocking=True)
            steps += 1
            samples += dense.shape[0]
        end.record()
        torch.cuda.current_stream().synchronize()
        res, loss = self._results(self._accum_host.tolist(), steps)
        timing = {"time_this_iter_s": time.perf_counter() - t0, "device_time_s": start.elapsed_time(end) / 1e3,
                  "samples_this_iter": samples, "steps_this_iter": steps, "engine": self.name,
                  "h2d_bytes_this_iter": loader.h2d_bytes - h2d0, "d2h_bytes_this_iter": 32 * steps}
        return res, loss, timing

    def evaluate_epoch(self, shard):
        loader = self._loader(shard)
        core = self.dlrm.core
        core.accum.zero_()
        steps = 0
        for dense, idx, y in loader:
            self.dlrm.eval_step(dense, idx, y)
            steps += 1
        return self._results(core.accum.tolist(), steps)

    def state_dict(self):
        torch.cuda.current_stream().synchronize()
        return self.dlrm.full_state_dict()

    def close(self):
        self.dlrm.close()
