"""Tracing / profiling helpers (the reference has none: SURVEY.md section 5.1).

* ``nvtx_range(name)``   -- NVTX range around ingest / forward / backward / reduce / step (shows up in
  Nsight tools); a no-op on hosts without CUDA.
* ``CudaStepTimer``      -- CUDA-event timing of K steps on the launching stream with a synchronise
  on both sides; ``max_over_ranks`` reduces with MAX so a multi-GPU number is the slowest rank.
* ``kernel_launch_csv``  -- parses an ``ncu --metrics gpu__time_duration.sum --csv`` log into
  ``(kernel, microseconds)`` rows (used to build the summaries under ``profiles/``).
"""
from __future__ import annotations

import contextlib
import csv
from typing import Iterator, List, Tuple


@contextlib.contextmanager
def nvtx_range(name: str) -> Iterator[None]:
    try:
        import torch
        enabled = torch.cuda.is_available()
    except ImportError:
        enabled = False
    if enabled:
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if enabled:
            torch.cuda.nvtx.range_pop()


class CudaStepTimer:
    def __init__(self):
        import torch
        self._torch = torch
        self.start = torch.cuda.Event(enable_timing=True)
        self.end = torch.cuda.Event(enable_timing=True)

    def __enter__(self):
        self._torch.cuda.synchronize()
        self.start.record()
        return self

    def __exit__(self, *exc):
        self.end.record()
        self._torch.cuda.synchronize()
        self.ms = self.start.elapsed_time(self.end)

    def max_over_ranks(self) -> float:
        import torch.distributed as dist
        if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
            return self.ms
        t = self._torch.tensor([self.ms], device="cuda", dtype=self._torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())


def kernel_launch_csv(path: str) -> List[Tuple[str, float]]:
    rows = [r for r in csv.reader(open(path)) if len(r) > 5]
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    h = rows[hdr]
    ki, vi, ui = h.index("Kernel Name"), h.index("Metric Value"), h.index("Metric Unit")
    out = []
    for r in rows[hdr + 1:]:
        try:
            v = float(r[vi].replace(",", ""))
        except ValueError:
            continue
        out.append((r[ki], v / 1000.0 if r[ui] in ("ns", "nsecond") else v))
    return out
