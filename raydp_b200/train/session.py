"""Per-worker training session: what ``train_func`` sees.

Parity: ``ray.train`` session calls in reference ``torch/estimator.py:226,248,252,263-270``:
``get_dataset_shard(name)``, ``report(metrics, checkpoint=)``,
``get_context().get_world_rank()``.
"""
from __future__ import annotations

import threading
from typing import Any, Dict, List, Optional

_local = threading.local()
_process_session: Optional["TrainSession"] = None


class TrainContext:
    def __init__(self, sess: "TrainSession"):
        self._s = sess

    def get_world_rank(self) -> int:
        return self._s.world_rank

    def get_world_size(self) -> int:
        return self._s.world_size

    def get_local_rank(self) -> int:
        return self._s.local_rank

    def get_local_world_size(self) -> int:
        return self._s.local_world_size

    def get_node_rank(self) -> int:
        return 0

    def get_trial_name(self) -> str:
        return self._s.run_name

    def get_storage(self):
        return self._s.storage_dir

    def get_device(self):
        return self._s.device


class TrainSession:
    def __init__(self, world_rank: int, world_size: int, local_rank: int, local_world_size: int,
                 dataset_shards: Dict[str, Any], device, run_name: str, storage_dir: str,
                 backend: str):
        self.world_rank = world_rank
        self.world_size = world_size
        self.local_rank = local_rank
        self.local_world_size = local_world_size
        self.dataset_shards = dataset_shards
        self.device = device
        self.run_name = run_name
        self.storage_dir = storage_dir
        self.backend = backend
        self.reports: List[Dict[str, Any]] = []

    def report(self, metrics: Dict[str, Any], checkpoint=None) -> None:
        entry: Dict[str, Any] = {"metrics": dict(metrics), "checkpoint": None}
        if checkpoint is not None and self.world_rank == 0:
            entry["checkpoint"] = checkpoint._persist(self.storage_dir).path
            self._prune_checkpoints()
        self.reports.append(entry)

    def _prune_checkpoints(self) -> None:
        """``CheckpointConfig(num_to_keep=k)``: the trainer leaves ``k`` in ``<storage>/.num_to_keep``; only the
        newest ``k`` persisted checkpoints stay on disk (the reference's XGBoost estimator keeps 1)."""
        import os
        import shutil
        marker = os.path.join(self.storage_dir, ".num_to_keep")
        if not os.path.exists(marker):
            return
        try:
            keep = int(open(marker, "r", encoding="utf-8").read().strip())
        except ValueError:
            return
        mine = [r["checkpoint"] for r in self.reports if r.get("checkpoint")]
        found = sorted((os.path.join(self.storage_dir, d) for d in os.listdir(self.storage_dir)
                        if d.startswith("checkpoint_")), key=os.path.getmtime)
        for path in found[:-keep] if keep > 0 else []:
            shutil.rmtree(path, ignore_errors=True)
            for r in self.reports:
                if r.get("checkpoint") == path:
                    r["checkpoint"] = None
        del mine


def _set_session(sess: Optional[TrainSession]) -> None:
    global _process_session
    _process_session = sess


def get_session() -> TrainSession:
    if _process_session is None:
        raise RuntimeError("not inside a training worker: no active train session")
    return _process_session


def get_context() -> TrainContext:
    return TrainContext(get_session())


def get_dataset_shard(name: Optional[str] = None):
    shards = get_session().dataset_shards
    if name is None:
        name = "train"
    return shards.get(name)


def report(metrics: Optional[Dict[str, Any]] = None, checkpoint=None, **kwargs) -> None:
    """``report({"loss": x}, checkpoint=...)``; the keyword form ``report(loss=x)`` of the older API is
    accepted too."""
    merged = dict(metrics or {})
    merged.update(kwargs)
    get_session().report(merged, checkpoint)


def get_checkpoint():
    """The checkpoint to resume from, or ``None``: ``TorchTrainer(resume_from_checkpoint=...)`` on the first
    attempt, the last checkpoint reported before a failure on a retry (``FailureConfig.max_failures``)."""
    import os
    from raydp_b200.train.checkpoint import Checkpoint
    marker = os.path.join(get_session().storage_dir, ".resume_from")
    if not os.path.exists(marker):
        return None
    with open(marker, "r", encoding="utf-8") as fh:
        path = fh.read().strip()
    return Checkpoint(path) if path and os.path.isdir(path) else None
