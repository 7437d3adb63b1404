"""In-worker torch helpers: ``prepare_model``, ``get_device``.

Parity: ``ray.train.torch.prepare_model`` as called by reference ``torch/estimator.py:243`` --
move the module to the worker's device and make it data parallel when world_size > 1.  On GPU
workers the wrapper is :class:`raydp_b200.parallel.FusedDataParallel` (hand-written NVLink
all-reduce fused with the optimizer) when the fused engine drives the step, otherwise stock
``DistributedDataParallel`` (the CPU/gloo path and the baseline).
"""
from __future__ import annotations

from raydp_b200.train.session import get_session


def get_device():
    return get_session().device


def prepare_model(model, move_to_device: bool = True, parallel_strategy: str = "ddp",
                  parallel_strategy_kwargs=None):
    import torch
    import torch.distributed as dist
    sess = get_session()
    if move_to_device:
        model = model.to(sess.device)
    if sess.world_size > 1 and parallel_strategy == "ddp" and dist.is_initialized():
        kwargs = dict(parallel_strategy_kwargs or {})
        if sess.device.type == "cuda":
            kwargs.setdefault("device_ids", [sess.device.index])
        model = torch.nn.parallel.DistributedDataParallel(model, **kwargs)
    return model


def unwrap_model(model):
    return model.module if hasattr(model, "module") and hasattr(model.module, "state_dict") \
        and type(model).__name__ in ("DistributedDataParallel", "FusedDataParallel") else model
