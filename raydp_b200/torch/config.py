"""Torch process-group configuration of the estimator.

Parity: reference ``python/raydp/torch/config.py:12-45`` defines a ``TorchConfig`` whose backend
hook imports Intel oneCCL bindings before process-group setup when ``use_ccl=True``.  This
image targets NVIDIA B200: ``backend="ccl"`` is honoured when the oneCCL bindings import, and
otherwise falls back to gloo with a warning instead of failing the run.
"""
from __future__ import annotations

import logging
from dataclasses import dataclass
from typing import Optional

from raydp_b200.train.config import TorchConfig as _BaseTorchConfig

logger = logging.getLogger(__name__)


def _ccl_importable() -> bool:
    for mod in ("oneccl_bindings_for_pytorch", "torch_ccl"):
        try:
            __import__(mod)
            return True
        except Exception:  # noqa: BLE001
            continue
    return False


@dataclass
class TorchConfig(_BaseTorchConfig):
    backend: Optional[str] = None

    def resolve_backend(self, use_gpu: bool) -> str:
        if self.backend == "ccl":
            if _ccl_importable():
                return "ccl"
            logger.warning("use_ccl=True but oneCCL bindings are not installed; using gloo")
            return "gloo"
        return super().resolve_backend(use_gpu)
