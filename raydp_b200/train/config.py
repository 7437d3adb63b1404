"""Trainer configuration objects.

Parity: the Ray Train / AIR config classes the reference passes to ``TorchTrainer``
(reference ``torch/estimator.py:331-347``): ``ScalingConfig(num_workers, resources_per_worker,
placement_strategy)``, ``RunConfig(failure_config=FailureConfig(max_failures))``,
``CheckpointConfig`` (reference ``xgboost/estimator.py:60-66``) and ``TorchConfig(backend)``
(reference ``torch/config.py:12-45``).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, Optional


@dataclass
class ScalingConfig:
    num_workers: int = 1
    use_gpu: Optional[bool] = None
    resources_per_worker: Optional[Dict[str, float]] = None
    placement_strategy: str = "PACK"
    trainer_resources: Optional[Dict[str, float]] = None

    def worker_resources(self) -> Dict[str, float]:
        res = {k: float(v) for k, v in (self.resources_per_worker or {}).items()}
        res.setdefault("CPU", 1.0)
        if self.use_gpu and "GPU" not in res:
            res["GPU"] = 1.0
        return res

    def wants_gpu(self) -> bool:
        if self.use_gpu is not None:
            return bool(self.use_gpu)
        return float((self.resources_per_worker or {}).get("GPU", 0)) > 0


@dataclass
class FailureConfig:
    max_failures: int = 0


@dataclass
class CheckpointConfig:
    num_to_keep: Optional[int] = None
    checkpoint_frequency: int = 0
    checkpoint_at_end: Optional[bool] = None


@dataclass
class RunConfig:
    name: Optional[str] = None
    storage_path: Optional[str] = None
    failure_config: FailureConfig = field(default_factory=FailureConfig)
    checkpoint_config: CheckpointConfig = field(default_factory=CheckpointConfig)
    verbose: int = 0


@dataclass
class TorchConfig:
    """``backend=None`` picks ``nccl`` on GPU workers and ``gloo`` on CPU workers."""
    backend: Optional[str] = None
    init_method: str = "env"
    timeout_s: int = 1800

    def resolve_backend(self, use_gpu: bool) -> str:
        if self.backend in (None, "auto"):
            return "nccl" if use_gpu else "gloo"
        return self.backend
