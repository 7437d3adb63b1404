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


def prepare_data_loader(data_loader, add_dist_sampler: bool = True, move_to_device: bool = True):
    """``ray.train.torch.prepare_data_loader``: shard a ``DataLoader`` over the worker group with a
    ``DistributedSampler`` (when it is not sharded already) and move every batch to the worker's device."""
    import torch
    from torch.utils.data import DataLoader, DistributedSampler
    sess = get_session()
    loader = data_loader
    if add_dist_sampler and sess.world_size > 1 and not isinstance(getattr(loader, "sampler", None), DistributedSampler):
        shuffle = isinstance(loader.sampler, torch.utils.data.RandomSampler)
        sampler = DistributedSampler(loader.dataset, num_replicas=sess.world_size, rank=sess.world_rank, shuffle=shuffle)
        loader = DataLoader(loader.dataset, batch_size=loader.batch_size, sampler=sampler,
                            num_workers=loader.num_workers, collate_fn=loader.collate_fn,
                            pin_memory=loader.pin_memory, drop_last=loader.drop_last)
    if not move_to_device or sess.device.type == "cpu":
        return loader

    class _OnDevice:
        def __init__(self, inner, device):
            self._inner, self._device = inner, device

        def __len__(self):
            return len(self._inner)

        def __iter__(self):
            def move(x):
                if torch.is_tensor(x):
                    return x.to(self._device, non_blocking=True)
                if isinstance(x, (list, tuple)):
                    return type(x)(move(v) for v in x)
                if isinstance(x, dict):
                    return {k: move(v) for k, v in x.items()}
                return x
            for batch in self._inner:
                yield move(batch)

        def __getattr__(self, name):
            return getattr(self._inner, name)

    return _OnDevice(loader, sess.device)
