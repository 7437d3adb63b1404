"""Data-parallel gradient reduction (SURVEY.md section 2.4 "DP", section 8 K1-K3).

``make_gradient_reducer`` picks the reduction used between backward and the optimizer:

* ``"p2p"`` (default on NVLink peers): hand-written one-shot / two-shot all-reduce over
  peer-mapped symmetric buffers (``raydp_b200.comm``), the B200-native replacement of the
  NCCL all-reduce issued by DDP in the reference (``prepare_model`` -> DDP, reference
  ``torch/estimator.py:243``);
* ``"nccl"``: ``torch.distributed.all_reduce`` -- the baseline and the fallback when peer access
  is unavailable.
"""
from __future__ import annotations

import logging
from typing import Any, Dict

import torch
import torch.distributed as dist

logger = logging.getLogger(__name__)


def __getattr__(name):
    if name == "FusedDataParallel":
        from raydp_b200.parallel.data_parallel import FusedDataParallel
        return FusedDataParallel
    raise AttributeError(name)


class NcclReducer:
    kind = "nccl"

    def __init__(self, grads: torch.Tensor):
        self.grads = grads

    def allreduce_(self) -> None:
        dist.all_reduce(self.grads)

    def close(self) -> None:
        pass


def make_gradient_reducer(grads: torch.Tensor, device: torch.device, extra: Dict[str, Any]):
    choice = extra.get("allreduce", "p2p")
    if choice == "p2p":
        try:
            from raydp_b200.comm import P2PReducer, p2p_available
            if p2p_available():
                return P2PReducer(grads, device)
            logger.warning("peer access between the ranks' GPUs is unavailable; using NCCL")
        except ImportError as exc:
            logger.warning("raydp_b200.comm is unavailable (%s); using NCCL", exc)
    return NcclReducer(grads)
