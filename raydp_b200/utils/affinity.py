"""CPU placement of a GPU worker process.

Eight trainer processes started by one launcher inherit one CPU mask; left alone, their Python
threads, the clock sampler and the copy-issuing code of the ingest pipeline pile onto the same few cores
and onto whichever NUMA node the scheduler likes, while each GPU's PCIe root hangs off ONE socket.
``bind_to_gpu`` narrows the calling process to the CPUs NVML reports as local to its GPU (intersected
with the mask the container allows) and, when several ranks share that set, to this rank's own slice of
it -- so pinned staging memory is first-touched and the copy calls are issued next to the GPU.
(The reference has no counterpart: it never places workers on GPUs at all)

Disabled with ``RAYDP_B200_BIND=0``.  Returns the CPU set chosen, or ``None`` when nothing was changed.
"""
from __future__ import annotations

import os
from typing import List, Optional


def _gpu_local_cpus(index: int) -> Optional[List[int]]:
    try:
        import pynvml
        pynvml.nvmlInit()
        try:
            handle = pynvml.nvmlDeviceGetHandleByIndex(index)
            words = pynvml.nvmlDeviceGetCpuAffinity(handle, (os.cpu_count() + 63) // 64)
        finally:
            pynvml.nvmlShutdown()
    except Exception:  # noqa: BLE001 - no NVML, no permission, MIG ...: leave the mask alone
        return None
    cpus = []
    for w, word in enumerate(words):
        for b in range(64):
            if (int(word) >> b) & 1:
                cpus.append(64 * w + b)
    return cpus or None


def plan_cpu_set(allowed: List[int], local: Optional[List[int]], rank_on_node: int, ranks_on_node: int,
                 min_cores: int = 2) -> List[int]:
    """Pure function (unit-tested): this rank's slice of the GPU-local CPUs it is allowed to use."""
    allowed = sorted(allowed)
    pool = sorted(set(allowed) & set(local)) if local else []
    if len(pool) < min_cores:
        pool = allowed
    ranks_on_node = max(1, ranks_on_node)
    per = max(min_cores, len(pool) // ranks_on_node)
    if per * ranks_on_node > len(pool):              # not enough cores for disjoint slices: overlap evenly
        start = (rank_on_node * max(1, len(pool) - per)) // max(1, ranks_on_node - 1) if ranks_on_node > 1 else 0
    else:
        start = rank_on_node * per
    return pool[start:start + per] or pool


def bind_to_gpu(gpu_index: int, rank_on_node: int, ranks_on_node: int) -> Optional[List[int]]:
    if os.environ.get("RAYDP_B200_BIND", "1") == "0" or not hasattr(os, "sched_setaffinity"):
        return None
    try:
        allowed = sorted(os.sched_getaffinity(0))
        # ranks whose GPUs report the SAME local CPUs share that pool (all of them on a single-socket box,
        # half of them on a two-socket one): slice the pool among exactly those ranks
        local = _gpu_local_cpus(gpu_index)
        if local is not None and ranks_on_node > 1:
            first = gpu_index - rank_on_node                   # GPU index of local rank 0
            pools = [_gpu_local_cpus(first + r) for r in range(ranks_on_node)]
            group = [r for r, p in enumerate(pools) if p == local]
            if rank_on_node in group:
                rank_on_node, ranks_on_node = group.index(rank_on_node), len(group)
        chosen = plan_cpu_set(allowed, local, rank_on_node, ranks_on_node)
        if chosen and set(chosen) != set(allowed):
            os.sched_setaffinity(0, chosen)
            return chosen
    except OSError:
        return None
    return None
