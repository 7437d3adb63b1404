"""Step engines behind ``TorchEstimator.train_func``.

``StockEngine`` restates the reference loop (reference ``torch/estimator.py:272-310``) with
stock parts only -- ``to_torch`` CPU batches, blocking ``.to(device)``, ``nn.Module`` forward,
``loss.item()`` every step, DDP all-reduce, ``optimizer.step()`` -- and is both the CPU/gloo
path and the GPU *baseline* (``engine="stock"``).

``FusedEngine`` (GPU workers, ``engine="auto"|"fused"``) keeps the same observable results but
moves the hot path onto hand-written sm_100a kernels: Arrow column buffers are DMA'd from
pinned shared memory on a side stream and packed by the ingest kernel, recognised layers run
on the tcgen05 GEMM / fused BN / fused loss kernels, the loss and metric sums stay on the
device (read once per epoch; the reported value is the same mean of per-batch means), and the
data-parallel gradient all-reduce is fused with the optimizer update over NVLink peer memory.
"""
from __future__ import annotations

import time
from typing import Dict, Tuple

import torch

from raydp_b200.train.torch_utils import prepare_model


def _call_model(model, inputs):
    return model(*inputs) if isinstance(inputs, (list, tuple)) else model(inputs)


def _to_device(obj, device, non_blocking: bool = False):
    if isinstance(obj, (list, tuple)):
        return [o.to(device, non_blocking=non_blocking) for o in obj]
    return obj.to(device, non_blocking=non_blocking)


def move_optimizer_state(optimizer, device) -> None:
    """Optimizer state created or loaded next to the CPU parameters (Adagrad's sums exist from
    construction on; Adam moments arrive with a pre-trained optimizer instance) has to follow the
    parameters to the device; 0-dim "step" counters stay host scalars, as torch.optim expects."""
    for st in optimizer.state.values():
        for k, v in list(st.items()):
            if torch.is_tensor(v) and v.dim() > 0 and v.device != device:
                st[k] = v.to(device)


class StockEngine:
    name = "stock"

    def __init__(self, config, sess, model, optimizer, criterion, scheduler, metrics):
        self.config = config
        self.sess = sess
        self.device = sess.device
        self.optimizer = optimizer
        self.criterion = criterion.to(self.device) if hasattr(criterion, "to") else criterion
        self.scheduler = scheduler
        self.metrics = metrics
        # The optimizer was built around the CPU parameters; .to(device) moves them in place
        # (Module.to keeps Parameter identity), so the optimizer keeps pointing at live tensors.
        self.model = prepare_model(model)
        move_optimizer_state(optimizer, self.device)
        self.autocast_dtype = config.get("compute_dtype") if self.device.type == "cuda" else None

    def _loader(self, shard):
        c = self.config
        return shard.to_torch(feature_columns=c["feature_columns"],
                              feature_column_dtypes=c["feature_types"],
                              label_column=c["label_column"], label_column_dtype=c["label_type"],
                              batch_size=c["batch_size"], drop_last=c["drop_last"],
                              null_fill=float(c.get("extra", {}).get("null_fill", float("nan"))))

    def train_epoch(self, shard) -> Tuple[Dict[str, float], float, Dict[str, float]]:
        self.model.train()
        t0 = time.perf_counter()
        total, steps, samples = 0.0, 0, 0
        for inputs, targets in self._loader(shard):
            inputs = _to_device(inputs, self.device)
            targets = targets.to(self.device)
            if self.autocast_dtype is not None:
                with torch.autocast("cuda", dtype=self.autocast_dtype):
                    outputs = _call_model(self.model, inputs)
                    loss = self.criterion(outputs.float(), targets)
            else:
                outputs = _call_model(self.model, inputs)
                loss = self.criterion(outputs, targets)
            total += loss.item()
            self.metrics.update(outputs.detach().float(), targets)
            samples += targets.size(0)
            self.optimizer.zero_grad()
            loss.backward()
            self.optimizer.step()
            if self.scheduler is not None:
                self.scheduler.step()
            steps += 1
        result = self.metrics.compute()
        self.metrics.reset()
        dt = time.perf_counter() - t0
        timing = {"time_this_iter_s": dt, "samples_this_iter": samples, "steps_this_iter": steps,
                  "engine": self.name}
        return result, total / max(1, steps), timing

    def evaluate_epoch(self, shard) -> Tuple[Dict[str, float], float]:
        self.model.eval()
        total, steps = 0.0, 0
        with torch.no_grad():
            for inputs, targets in self._loader(shard):
                inputs = _to_device(inputs, self.device)
                targets = targets.to(self.device)
                outputs = _call_model(self.model, inputs)
                total += self.criterion(outputs, targets).item()
                self.metrics.update(outputs.float(), targets)
                steps += 1
        result = self.metrics.compute()
        self.metrics.reset()
        return result, total / max(1, steps)

    def state_dict(self) -> Dict[str, torch.Tensor]:
        module = self.model.module if hasattr(self.model, "module") else self.model
        return {k: v.detach().cpu() for k, v in module.state_dict().items()}

    def close(self) -> None:
        pass


def make_engine(config, sess, model, optimizer, criterion, scheduler, metrics):
    """Pick the step engine.  CUDA workers never run the stock loop unless ``engine="stock"`` asks for
    the baseline: a module the lowering recognises (``raydp_b200.torch.lowering``: the reference's own
    ``NYC_Model`` / notebook ``Dlrm`` classes included) runs on the whole-step fused engines, anything
    else on :class:`GenericGpuEngine` (tcgen05 Linear layers + fused NVLink data parallelism + GPU ingest)."""
    choice = config.get("engine", "auto")
    if choice not in ("auto", "fused", "generic", "stock"):
        raise ValueError(f"engine={choice!r}: expected auto | fused | generic | stock")
    if sess.device.type == "cuda" and choice != "stock":
        from raydp_b200.torch.fused_engine import DLRMEngine, FusedEngine, select_fused
        from raydp_b200.torch.generic_engine import GenericGpuEngine
        plan, why = (None, "engine='generic' was requested") if choice == "generic" else \
            select_fused(config, model, optimizer, criterion)
        if plan is not None and plan.kind == "dlrm":
            return DLRMEngine(config, sess, model, optimizer, criterion, scheduler, metrics, plan)
        if plan is not None:
            return FusedEngine(config, sess, model, optimizer, criterion, scheduler, metrics, plan)
        if choice == "fused":
            raise RuntimeError(f"engine='fused' was requested but is not applicable: {why}")
        return GenericGpuEngine(config, sess, model, optimizer, criterion, scheduler, metrics, why)
    if choice in ("fused", "generic"):
        raise RuntimeError(f"engine={choice!r} needs CUDA workers (resources_per_worker={{'GPU': 1}})")
    return StockEngine(config, sess, model, optimizer, criterion, scheduler, metrics)
