"""Tracing / profiling helpers (the reference has none: SURVEY.md section 5.1).

* ``nvtx_range(name)``   -- NVTX range around ingest / forward / backward / reduce / step (shows up in
  Nsight tools); a no-op on hosts without CUDA.
* ``ClockSampler``       -- samples SM clocks and throttle reasons with ``nvidia-smi`` DURING a timed region
  (every benchmark record carries its own ``clocks``).
* ``CudaStepTimer``      -- CUDA-event timing of K steps on the launching stream with a synchronise
  on both sides; ``max_over_ranks`` reduces with MAX so a multi-GPU number is the slowest rank.
* ``kernel_launch_csv``  -- parses an ``ncu --metrics gpu__time_duration.sum --csv`` log into
  ``(kernel, microseconds)`` rows (used to build the summaries under ``profiles/``).
"""
from __future__ import annotations

import contextlib
import csv
import subprocess
import threading
import time
from typing import Iterator, List, Tuple


_nvtx_enabled = None


def _nvtx_on() -> bool:
    """NVTX is on for CUDA processes unless RAYDP_B200_NVTX=0 (decided once per process)."""
    global _nvtx_enabled
    if _nvtx_enabled is None:
        import os
        try:
            import torch
            _nvtx_enabled = torch.cuda.is_available() and os.environ.get("RAYDP_B200_NVTX", "1") != "0"
        except ImportError:
            _nvtx_enabled = False
    return _nvtx_enabled


@contextlib.contextmanager
def nvtx_range(name: str) -> Iterator[None]:
    enabled = _nvtx_on()
    if enabled:
        import torch
        torch.cuda.nvtx.range_push(name)
    try:
        yield
    finally:
        if enabled:
            torch.cuda.nvtx.range_pop()


class ClockSampler:
    """Samples SM clocks / throttle reasons of this rank's GPU during the timed region."""

    QUERY = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,"
             "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
             "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.QUERY}",
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._read, daemon=True)
            self.thread.start()
        except OSError:
            self.proc = None

    def _read(self):
        for line in self.proc.stdout:
            self.lines.append((time.time(), line.strip()))

    def wait_first_sample(self, timeout: float = 3.0) -> None:
        t0 = time.time()
        while not self.lines and time.time() - t0 < timeout:
            time.sleep(0.01)

    def stop(self, t_begin: float = 0.0, t_end: float = float("inf")) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.06)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        inside = [l for ts, l in self.lines if t_begin <= ts <= t_end + 0.03]
        window = "timed region"
        if not inside:      # region shorter than the sampling period: use the samples taken under load
            inside = [l for ts, l in self.lines if ts >= t_begin - 0.25]
            window = "warm-up + timed region (timed region shorter than the 20 ms sampling period)"
        for line in inside:
            parts = [p.strip() for p in line.split(",")]
            if len(parts) < 6:
                continue
            try:
                sm.append(float(parts[0]))
                mx.append(float(parts[1]))
            except ValueError:
                continue
            for name, val in zip(names, parts[2:6]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "window": window, "reasons": sorted(reasons)}


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
