"""Training metrics by name.

Parity: reference ``python/raydp/torch/torch_metrics.py:21-55`` wraps ``torchmetrics``: a list of
metric *names* (with optional kwargs from ``metrics_config[name]``) or metric instances;
``update / compute / reset``; ``compute()`` returns ``{name: float}``.  ``torchmetrics`` is used
when importable; otherwise the built-in running-sum metrics below (device-resident
accumulators, one ``.item()`` per ``compute``) cover the names the reference's examples use.
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Union

import torch


class _Running:
    """Base: accumulate sums on the device of the first update; sync to host only in compute."""

    def __init__(self, **_kw):
        self._acc: Optional[torch.Tensor] = None

    def _add(self, values: List[torch.Tensor]) -> None:
        vec = torch.stack([v.detach().to(torch.float64).reshape(()) for v in values])
        self._acc = vec if self._acc is None else self._acc + vec

    def add_sums(self, sums: torch.Tensor) -> None:
        """Fold externally computed partial sums (the fused loss kernel produces them)."""
        sums = sums.detach().to(torch.float64)
        self._acc = sums if self._acc is None else self._acc + sums

    def reset(self) -> None:
        self._acc = None

    def to(self, device):
        if self._acc is not None:
            self._acc = self._acc.to(device)
        return self


class MeanAbsoluteError(_Running):
    def update(self, preds, target):
        self._add([(preds.float() - target.float()).abs().sum(),
                   torch.tensor(float(target.numel()), device=preds.device)])

    def compute(self):
        return (self._acc[0] / self._acc[1]).float() if self._acc is not None else torch.tensor(float("nan"))


class MeanSquaredError(_Running):
    def __init__(self, squared: bool = True, **kw):
        super().__init__(**kw)
        self.squared = squared

    def update(self, preds, target):
        self._add([((preds.float() - target.float()) ** 2).sum(),
                   torch.tensor(float(target.numel()), device=preds.device)])

    def compute(self):
        if self._acc is None:
            return torch.tensor(float("nan"))
        mse = self._acc[0] / self._acc[1]
        return (mse if self.squared else mse.sqrt()).float()


class Accuracy(_Running):
    """Binary accuracy on logits/probabilities (threshold) or argmax over a class dimension."""

    def __init__(self, threshold: float = 0.5, task: str = "binary", logits: Optional[bool] = None, **kw):
        super().__init__(**kw)
        self.threshold = threshold
        self.logits = logits

    def update(self, preds, target):
        if preds.dim() > 1 and preds.shape[-1] > 1:
            hard = preds.argmax(dim=-1)
        else:
            p = preds.reshape(-1).float()
            as_logits = self.logits if self.logits is not None else bool((p.min() < 0) | (p.max() > 1))
            hard = (torch.sigmoid(p) if as_logits else p) >= self.threshold
        correct = (hard.reshape(-1).long() == target.reshape(-1).long()).sum()
        self._add([correct, torch.tensor(float(target.numel()), device=preds.device)])

    def compute(self):
        return (self._acc[0] / self._acc[1]).float() if self._acc is not None else torch.tensor(float("nan"))


class R2Score(_Running):
    def update(self, preds, target):
        p, t = preds.float().reshape(-1), target.float().reshape(-1)
        self._add([((p - t) ** 2).sum(), t.sum(), (t ** 2).sum(),
                   torch.tensor(float(t.numel()), device=preds.device)])

    def compute(self):
        if self._acc is None:
            return torch.tensor(float("nan"))
        sse, s, sq, n = self._acc
        return (1 - sse / (sq - s * s / n)).float()


_BUILTIN = {"MeanAbsoluteError": MeanAbsoluteError, "MeanSquaredError": MeanSquaredError,
            "Accuracy": Accuracy, "BinaryAccuracy": Accuracy, "R2Score": R2Score}


def _lookup(name: str):
    try:
        import torchmetrics  # type: ignore
        if hasattr(torchmetrics, name):
            return getattr(torchmetrics, name)
    except ImportError:
        pass
    if name in _BUILTIN:
        return _BUILTIN[name]
    raise ValueError(f"unknown metric {name!r}; built-in metrics: {sorted(_BUILTIN)} "
                     "(install torchmetrics for the full catalogue)")


class TorchMetric:
    def __init__(self, metrics_name: Optional[List[Union[str, Any]]],
                 metrics_config: Optional[Dict[str, Dict[str, Any]]]):
        self._metrics_name = metrics_name
        self._metrics_func: Dict[str, Any] = {}
        if metrics_name is not None:
            assert isinstance(metrics_name, list), "metrics_name must be a list"
            for metric in metrics_name:
                if isinstance(metric, str):
                    kwargs = (metrics_config or {}).get(metric) or {}
                    self._metrics_func[metric] = _lookup(metric)(**kwargs)
                elif all(hasattr(metric, a) for a in ("update", "compute", "reset")):
                    self._metrics_func[metric.__class__.__name__] = metric
                else:
                    raise Exception("Unsupported parameter, we only support list of metric names "
                                    "or metric instances with update/compute/reset")

    @property
    def names(self) -> List[str]:
        return list(self._metrics_func)

    def get(self, name: str):
        return self._metrics_func.get(name)

    def to(self, device):
        for m in self._metrics_func.values():
            if hasattr(m, "to"):
                m.to(device)
        return self

    def update(self, preds, targets) -> None:
        for metric in self._metrics_func.values():
            metric.update(preds, targets)

    def compute(self) -> Dict[str, float]:
        return {name: float(metric.compute().item() if hasattr(metric.compute(), "item")
                            else metric.compute())
                for name, metric in self._metrics_func.items()}

    def reset(self) -> None:
        for metric in self._metrics_func.values():
            metric.reset()
