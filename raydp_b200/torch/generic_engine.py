"""GPU engine for ARBITRARY user modules (the fallback behind the fully fused engines).

The reference accepts any ``nn.Module`` or creator function and wraps it in DDP
(``python/raydp/torch/estimator.py:180-187,243``), feeds it from ``dataset.to_torch`` (``:226-232``)
and reads ``loss.item()`` every step (``:280``).  A module that :mod:`raydp_b200.torch.lowering` cannot
map onto the whole-step fused kernels still gets the B200 path, piece by piece:

* every plain ``nn.Linear`` is replaced by :class:`raydp_b200.ops.linear.FusedLinear` adopting the
  same Parameter objects (tcgen05 forward / dgrad / wgrad; odd shapes run zero-padded), with
  ``Linear -> ReLU`` pairs of a ``Sequential`` fused into one epilogue; everything else in the module
  runs as the user wrote it, under bf16 autocast;
* data parallelism is :class:`raydp_b200.parallel.FusedDataParallel`: flat gradient buckets in
  registered symmetric memory, NVLink/NVLS all-reduce fused with Adam / AdamW / SGD and overlapped
  with backward -- never stock DDP + NCCL.  Other optimizers keep their own ``step()`` after the fused
  gradient averaging;
* batches come from :class:`raydp_b200.ops.ingest.ArrowGpuLoader` (pinned Arrow buffers -> DMA ->
  ingest kernel): one fp32 ``[B, F]`` tensor for a single feature dtype, or -- with a list of
  ``feature_types`` -- one ``[B, 1]`` tensor per column passed as ``model(*inputs)`` (reference
  docstring ``:115-121``, the DLRM notebook's calling convention);
* loss and metric sums stay on the device and are read once per epoch; the reported loss is the
  same mean of per-batch means.
"""
from __future__ import annotations

import time
from typing import Dict, List, Tuple

import torch

from raydp_b200 import ops
from raydp_b200.ops.ingest import ArrowGpuLoader
from raydp_b200.ops.linear import swap_linears
from raydp_b200.utils.profiling import nvtx_range

_INT_DTYPES = (torch.int64, torch.int32, torch.int16, torch.int8, torch.uint8, torch.bool)


class GenericGpuEngine:
    name = "fused-generic"

    def __init__(self, config, sess, model, optimizer, criterion, scheduler, metrics, why: str = ""):
        from raydp_b200.parallel.data_parallel import FusedDataParallel
        self.config, self.sess, self.device = config, sess, sess.device
        self.optimizer, self.scheduler, self.metrics = optimizer, scheduler, metrics
        self.criterion = criterion.to(self.device) if hasattr(criterion, "to") else criterion
        self.batch = int(config["batch_size"])
        self.world = sess.world_size
        self.why_not_fused = why
        if not ops.kernels_available():
            raise RuntimeError("CUDA worker without librdp_kernels.so: run __graft_entry__.build() first")
        model = model.to(self.device)             # keeps Parameter identity: the optimizer stays attached
        from raydp_b200.torch.engine import move_optimizer_state
        move_optimizer_state(optimizer, self.device)
        self.n_fused_linears = swap_linears(model)
        for m in model.modules():                 # gradients live in dense flat buckets (peer-reducible)
            if isinstance(m, (torch.nn.Embedding, torch.nn.EmbeddingBag)) and m.sparse:
                m.sparse = False
        self.model = model
        self.dp = FusedDataParallel(model, optimizer, self.device,
                                    bucket_mb=int(config.get("extra", {}).get("bucket_mb", 512)))
        types = config.get("feature_types")
        self.per_column = isinstance(types, (list, tuple))
        cols = list(config["feature_columns"])
        if self.per_column:
            if len(types) != len(cols):
                raise ValueError("feature_types must have one dtype per feature column")
            self.float_cols = [c for c, t in zip(cols, types) if t not in _INT_DTYPES]
            self.int_cols = [c for c, t in zip(cols, types) if t in _INT_DTYPES]
            self.col_plan = []                                   # (is_int, position, dtype) per feature column
            fi = ii = 0
            for c, t in zip(cols, types):
                if t in _INT_DTYPES:
                    self.col_plan.append((True, ii, t))
                    ii += 1
                else:
                    self.col_plan.append((False, fi, t))
                    fi += 1
        else:
            self.float_cols, self.int_cols, self.col_plan = cols, [], []
        self.feature_dtype = types if (types is not None and not self.per_column) else torch.float32
        self.label_dtype = config.get("label_type") or torch.float32
        self.autocast = config.get("compute_dtype") or torch.bfloat16
        self._loss_sum = torch.zeros((), dtype=torch.float32, device=self.device)
        self._loaders: Dict[int, ArrowGpuLoader] = {}
        self.d2h_bytes = 0

    # ------------------------------------------------------------------ data
    def _loader(self, shard) -> ArrowGpuLoader:
        key = id(shard)
        if key not in self._loaders:
            c = self.config
            self._loaders[key] = ArrowGpuLoader(
                shard, self.float_cols, c["label_column"], self.batch, c["drop_last"], self.device,
                out_dtype=torch.float32, sparse_columns=self.int_cols or None,
                table_rows=[0] * len(self.int_cols) if self.int_cols else None,
                null_fill=float(c.get("extra", {}).get("null_fill", float("nan"))))
            self._loaders[key]._prepare()
        return self._loaders[key]

    def _inputs(self, item):
        """Loader item -> (model inputs, target [B, 1])."""
        if self.int_cols:
            x, s, y = item
        else:
            (x, y), s = item, None
        nf = self._loaders[next(iter(self._loaders))].n_features if self._loaders else len(self.float_cols)
        if self.per_column:
            inputs: List[torch.Tensor] = []
            for is_int, pos, dt in self.col_plan:
                if is_int:
                    inputs.append(s[:, pos:pos + 1].to(dt))
                else:
                    col = x[:, pos:pos + 1]
                    inputs.append(col if dt in (None, torch.float32) else col.to(dt))
        else:
            inputs = x[:, :nf]
            if self.feature_dtype not in (None, torch.float32):
                inputs = inputs.to(self.feature_dtype)
        target = y.unsqueeze(1)
        if self.label_dtype not in (None, torch.float32):
            target = target.to(self.label_dtype)
        return inputs, target

    def _forward(self, inputs):
        with torch.autocast("cuda", dtype=self.autocast):
            out = self.model(*inputs) if isinstance(inputs, list) else self.model(inputs)
        return out.float() if out.is_floating_point() else out

    # ------------------------------------------------------------------ epochs
    def train_epoch(self, shard) -> Tuple[Dict[str, float], float, Dict[str, float]]:
        self.model.train()
        loader = self._loader(shard)
        self._loss_sum.zero_()
        start, end = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0 = time.perf_counter()
        start.record()
        steps = samples = 0
        h2d0 = loader.h2d_bytes
        launches0 = ops.launch_count()
        for item in loader:
            with nvtx_range("rdp.generic.step"):
                inputs, target = self._inputs(item)
                out = self._forward(inputs)
                loss = self.criterion(out, target)
                self._loss_sum += loss.detach()
                self.metrics.update(out.detach(), target)
                self.dp.backward_and_step(loss)
                if self.scheduler is not None:
                    self.scheduler.step()
            steps += 1
            samples += target.shape[0]
        end.record()
        torch.cuda.current_stream().synchronize()
        if self.world > 1:
            from raydp_b200.comm import check_status
            check_status()
        dt = time.perf_counter() - t0
        total = float(self._loss_sum.item())          # the ONE device->host read of the epoch
        result = self.metrics.compute()
        self.metrics.reset()
        timing = {"time_this_iter_s": dt, "device_time_s": start.elapsed_time(end) / 1e3,
                  "samples_this_iter": samples, "steps_this_iter": steps, "engine": self.name,
                  "h2d_bytes_this_iter": loader.h2d_bytes - h2d0, "d2h_bytes_this_iter": 4,
                  "fused_linears": self.n_fused_linears, "kernel_launches": ops.launch_count() - launches0}
        return result, total / max(1, steps), timing

    def evaluate_epoch(self, shard) -> Tuple[Dict[str, float], float]:
        self.model.eval()
        loader = self._loader(shard)
        self._loss_sum.zero_()
        steps = 0
        with torch.no_grad():
            for item in loader:
                inputs, target = self._inputs(item)
                out = self._forward(inputs)
                self._loss_sum += self.criterion(out, target)
                self.metrics.update(out, target)
                steps += 1
        total = float(self._loss_sum.item())
        result = self.metrics.compute()
        self.metrics.reset()
        return result, total / max(1, steps)

    def state_dict(self) -> Dict[str, torch.Tensor]:
        torch.cuda.current_stream().synchronize()
        return {k: v.detach().cpu().clone() for k, v in self.model.state_dict().items()}

    def close(self) -> None:
        self.dp.close()
