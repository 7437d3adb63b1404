"""The ``horovod.torch`` names the reference's example uses (``examples/horovod_nyctaxi.py:6-8,120-138``), on top of
the worker group's ``torch.distributed`` process group, plus ``ray.train.horovod.HorovodTrainer``.

``raydp_b200.compat`` registers this module as ``horovod.torch`` (and ``HorovodTrainer`` under ``ray.train.horovod``)
when Horovod is absent: ``hvd.init()`` is a no-op (the trainer has formed the group), ``hvd.DistributedOptimizer``
averages the gradients over the group in ``step()`` -- one flat all-reduce per step -- before the wrapped optimizer
runs.  On GPU workers prefer :class:`raydp_b200.parallel.FusedDataParallel` (peer-memory all-reduce fused with the
optimizer); this module exists so that scripts written against Horovod run as they are.
"""
from __future__ import annotations

from typing import Any, Dict, Iterable, Optional, Tuple

Average, Sum = "average", "sum"


def _dist():
    import torch.distributed as dist
    return dist


def init(*_args, **_kwargs) -> None:
    """The process group exists already when a trainer runs the loop; a single process needs none."""


def shutdown() -> None:
    pass


def is_initialized() -> bool:
    return True


def size() -> int:
    dist = _dist()
    return dist.get_world_size() if dist.is_initialized() else 1


def rank() -> int:
    dist = _dist()
    return dist.get_rank() if dist.is_initialized() else 0


def local_rank() -> int:
    try:
        from raydp_b200.train.session import get_context
        return get_context().get_local_rank()
    except Exception:  # noqa: BLE001 - outside a training session
        return 0


def local_size() -> int:
    try:
        from raydp_b200.train.session import get_context
        return get_context().get_local_world_size()
    except Exception:  # noqa: BLE001
        return 1


def allreduce(tensor, average: Optional[bool] = None, name: Optional[str] = None, op: Optional[str] = None):
    import torch
    dist = _dist()
    out = tensor.clone()
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(out)
        mean = (op == Average) if op is not None else (average is None or average)
        if mean:
            out = out / dist.get_world_size() if out.is_floating_point() else torch.div(out, dist.get_world_size(),
                                                                                        rounding_mode="floor")
    return out


def broadcast(tensor, root_rank: int = 0, name: Optional[str] = None):
    dist = _dist()
    out = tensor.clone()
    if dist.is_initialized() and dist.get_world_size() > 1:
        dist.broadcast(out, src=root_rank)
    return out


def broadcast_parameters(params, root_rank: int = 0) -> None:
    """``params``: a ``state_dict()`` or an iterable of ``(name, tensor)``; updated in place from ``root_rank``."""
    import torch
    dist = _dist()
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return
    items: Iterable[Tuple[str, Any]] = params.items() if isinstance(params, dict) else params
    for _, t in items:
        if torch.is_tensor(t):
            dist.broadcast(t.data if hasattr(t, "data") else t, src=root_rank)


def broadcast_optimizer_state(optimizer, root_rank: int = 0) -> None:
    dist = _dist()
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return
    box = [optimizer.state_dict() if dist.get_rank() == root_rank else None]
    dist.broadcast_object_list(box, src=root_rank)
    if dist.get_rank() != root_rank:
        optimizer.load_state_dict(box[0])


def broadcast_object(obj, root_rank: int = 0, name: Optional[str] = None):
    dist = _dist()
    if not (dist.is_initialized() and dist.get_world_size() > 1):
        return obj
    box = [obj if dist.get_rank() == root_rank else None]
    dist.broadcast_object_list(box, src=root_rank)
    return box[0]


class _DistributedOptimizer:
    """Wraps a ``torch.optim.Optimizer``: ``step()`` first reduces every gradient over the group (one all-reduce
    of a flat buffer per dtype), then runs the wrapped optimizer."""

    def __init__(self, optimizer, named_parameters=None, op: str = Average, **_ignored):
        self._opt = optimizer
        self._op = op
        if named_parameters is not None:
            self._params = [p for _, p in named_parameters]
        else:
            self._params = [p for g in optimizer.param_groups for p in g["params"]]

    def __getattr__(self, name):
        return getattr(self._opt, name)

    def synchronize(self) -> None:
        import torch
        dist = _dist()
        if not (dist.is_initialized() and dist.get_world_size() > 1):
            return
        by_type: Dict[Any, list] = {}
        for p in self._params:
            if p.grad is not None:
                by_type.setdefault((p.grad.dtype, p.grad.device), []).append(p.grad)
        for grads in by_type.values():
            flat = torch.cat([g.reshape(-1) for g in grads])
            dist.all_reduce(flat)
            if self._op == Average:
                flat /= dist.get_world_size()
            off = 0
            for g in grads:
                g.copy_(flat[off:off + g.numel()].view_as(g))
                off += g.numel()

    def step(self, closure=None):
        self.synchronize()
        return self._opt.step(closure) if closure is not None else self._opt.step()

    def zero_grad(self, *args, **kwargs):
        return self._opt.zero_grad(*args, **kwargs)


def DistributedOptimizer(optimizer, named_parameters=None, op: str = Average, **kwargs):  # noqa: N802 (Horovod's name)
    return _DistributedOptimizer(optimizer, named_parameters=named_parameters, op=op, **kwargs)


def horovod_trainer_class():
    """``ray.train.horovod.HorovodTrainer``: the worker-group trainer under Horovod's name (gloo / NCCL group)."""
    from raydp_b200.train.trainer import TorchTrainer

    class HorovodTrainer(TorchTrainer):
        def __init__(self, train_loop_per_worker, *, train_loop_config=None, horovod_config=None, **kwargs):
            super().__init__(train_loop_per_worker, train_loop_config=train_loop_config, **kwargs)

    return HorovodTrainer
