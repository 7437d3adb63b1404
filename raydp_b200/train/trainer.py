"""``TorchTrainer``: run a training function on N worker processes, one GPU (or CPU) each.

Parity: Ray Train's ``TorchTrainer(train_loop_per_worker, train_loop_config, scaling_config,
run_config, torch_config, datasets).fit() -> result.checkpoint`` exactly as the reference
drives it (reference ``torch/estimator.py:348-356``).  Semantics kept:
datasets are split into equal per-worker shards, ``FailureConfig(max_failures)`` restarts the
whole worker group, rank 0's reports/checkpoint form the result.

Two launch modes:

* **actors** (default): N ``TrainWorker`` actors inside a placement group with the requested
  strategy; each gets one GPU through the runtime's ``GPU`` resource.
* **SPMD**: when the calling process is itself one rank of a ``torchrun`` world whose size
  equals ``num_workers`` the training function runs in-process on that rank (one process per
  GPU is already given), and ``fit()`` is a collective call made by every rank.
"""
from __future__ import annotations

import os
import tempfile
import time
import uuid
from typing import Any, Callable, Dict, List, Optional

from raydp_b200 import runtime
from raydp_b200.train.checkpoint import Checkpoint
from raydp_b200.train.config import RunConfig, ScalingConfig, TorchConfig
from raydp_b200.train.session import TrainSession, _set_session
from raydp_b200.utils import find_free_port


class TrainingFailedError(RuntimeError):
    pass


class Result:
    def __init__(self, metrics: Dict[str, Any], checkpoint: Optional[Checkpoint],
                 metrics_history: List[Dict[str, Any]], path: str, error: Optional[BaseException] = None):
        self.metrics = metrics
        self.checkpoint = checkpoint
        self.metrics_history = metrics_history
        self.path = path
        self.error = error

    @property
    def metrics_dataframe(self):
        import pandas as pd
        return pd.DataFrame(self.metrics_history)

    def __repr__(self):
        return f"Result(metrics={self.metrics}, checkpoint={self.checkpoint})"


def _spmd_world() -> Optional[int]:
    if os.environ.get("RAYDP_B200_SPMD", "1") == "0":
        return None
    if "RANK" in os.environ and "WORLD_SIZE" in os.environ and not os.environ.get("RAYDP_B200_ACTOR_ID"):
        return int(os.environ["WORLD_SIZE"])
    return None


def _worker_log_tail(worker, limit: int = 2000) -> str:
    try:
        logs = runtime.worker_logs(str(getattr(worker, "_actor_id", ""))[:8], tail_bytes=limit)
        for rec in logs.values():
            lines = [ln for ln in rec["text"].splitlines() if not ln.startswith((":job_id:", ":actor_name:"))]
            if lines:
                return "\n--- last output of the worker (" + rec["tag"] + ") ---\n" + "\n".join(lines[-25:])
    except Exception:  # noqa: BLE001 - diagnostics must not mask the failure
        pass
    return ""


def run_worker(train_func: Callable, config: Optional[dict], rank: int, world: int,
               local_rank: int, master_addr: str, master_port: int, backend: str, use_gpu: bool,
               shards: Dict[str, Any], run_name: str, storage_dir: str,
               init_pg: bool = True) -> List[Dict[str, Any]]:
    """Body of one training rank (shared by actor workers and SPMD mode)."""
    import torch
    import torch.distributed as dist
    device = torch.device("cpu")
    if use_gpu:
        if not torch.cuda.is_available():
            raise RuntimeError("a GPU worker was requested but CUDA is not available in this process")
        index = local_rank if torch.cuda.device_count() > local_rank else 0
        torch.cuda.set_device(index)
        device = torch.device("cuda", index)
        from raydp_b200.utils.affinity import bind_to_gpu
        bind_to_gpu(index, local_rank if local_rank < world else rank % max(1, world), world)
    created_pg = False
    if init_pg and world > 1 and not dist.is_initialized():
        # A driver that itself runs under torchrun hands its environment down to the actors: with
        # TORCHELASTIC_USE_AGENT_STORE set, env:// rendezvous makes EVERY rank a client of the launcher's store at
        # MASTER_ADDR:MASTER_PORT -- ours is a fresh port nobody serves, so the group would wait out the store
        # timeout.  This group is ours: rank 0 serves the store.
        for var in ("TORCHELASTIC_USE_AGENT_STORE", "TORCHELASTIC_RUN_ID", "TORCHELASTIC_RESTART_COUNT",
                    "TORCHELASTIC_MAX_RESTARTS", "TORCHELASTIC_ERROR_FILE", "GROUP_RANK", "ROLE_RANK",
                    "ROLE_WORLD_SIZE", "GROUP_WORLD_SIZE", "LOCAL_WORLD_SIZE"):
            os.environ.pop(var, None)
        os.environ["MASTER_ADDR"] = master_addr
        os.environ["MASTER_PORT"] = str(master_port)
        os.environ["RANK"], os.environ["WORLD_SIZE"] = str(rank), str(world)
        os.environ["LOCAL_RANK"] = str(local_rank)
        kwargs = {}
        if backend == "nccl":
            kwargs["device_id"] = device
        pg_timeout = os.environ.get("RAYDP_B200_PG_TIMEOUT_S")
        if pg_timeout:
            import datetime
            kwargs["timeout"] = datetime.timedelta(seconds=float(pg_timeout))
        dist.init_process_group(backend=backend, rank=rank, world_size=world, **kwargs)
        created_pg = True
    sess = TrainSession(rank, world, local_rank, world, shards, device, run_name, storage_dir, backend)
    _set_session(sess)
    try:
        if config is None:
            train_func()
        else:
            train_func(config)
        return sess.reports
    finally:
        _set_session(None)
        if created_pg:
            try:
                dist.destroy_process_group()
            except Exception:  # noqa: BLE001
                pass


class TrainWorker:
    """Actor hosting one rank."""

    def __init__(self, rank: int, world: int):
        self.rank, self.world = rank, world

    def node_ip(self) -> str:
        return runtime.get_node_ip_address()

    def run(self, train_func, config, master_addr, master_port, backend, use_gpu, shards,
            run_name, storage_dir):
        local_rank = self.rank
        if use_gpu:
            gpus = runtime.get_runtime_context().get_assigned_gpus()
            local_rank = int(gpus[0]) if gpus else self.rank
        return run_worker(train_func, config, self.rank, self.world, local_rank,
                          master_addr, master_port, backend, use_gpu, shards, run_name, storage_dir)

    def shutdown(self):
        runtime.exit_actor()


class TorchTrainer:
    def __init__(self, train_loop_per_worker: Callable, *, train_loop_config: Optional[dict] = None,
                 scaling_config: Optional[ScalingConfig] = None, run_config: Optional[RunConfig] = None,
                 torch_config: Optional[TorchConfig] = None, datasets: Optional[Dict[str, Any]] = None,
                 dataset_config=None, resume_from_checkpoint=None, metadata=None):
        self._fn = train_loop_per_worker
        self._config = train_loop_config
        self._scaling = scaling_config or ScalingConfig()
        self._run = run_config or RunConfig()
        self._torch = torch_config or TorchConfig()
        self._datasets = {k: v for k, v in (datasets or {}).items() if v is not None}
        self._resume = resume_from_checkpoint

    # ------------------------------------------------------------------ shards
    def _shards_for(self, world: int) -> List[Dict[str, Any]]:
        out: List[Dict[str, Any]] = [dict() for _ in range(world)]
        for name, ds in self._datasets.items():
            parts = ds.split(world, equal=True) if world > 1 else [ds]
            for r in range(world):
                out[r][name] = parts[r]
        return out

    def _storage(self) -> str:
        base = self._run.storage_path or os.path.join(tempfile.gettempdir(), "raydp_b200_results")
        name = self._run.name or f"TorchTrainer_{time.strftime('%Y%m%d_%H%M%S')}_{uuid.uuid4().hex[:6]}"
        path = os.path.join(base, name)
        os.makedirs(path, exist_ok=True)
        return path

    # ------------------------------------------------------------------ fit
    def fit(self) -> Result:
        storage = self._storage()
        run_name = os.path.basename(storage)
        keep = self._run.checkpoint_config.num_to_keep
        if keep:
            with open(os.path.join(storage, ".num_to_keep"), "w", encoding="utf-8") as fh:
                fh.write(str(int(keep)))
        max_failures = self._run.failure_config.max_failures
        attempt, last_error = 0, None
        while True:
            self._publish_resume_point(storage, attempt, self._resume)
            try:
                spmd = _spmd_world()
                if spmd is not None and spmd == self._scaling.num_workers:
                    reports = self._fit_spmd(run_name, storage)
                else:
                    reports = self._fit_actors(run_name, storage)
                return self._result(reports, storage)
            except (runtime.ActorDiedError, runtime.TaskError, TrainingFailedError) as exc:
                last_error = exc
                attempt += 1
                if max_failures >= 0 and attempt > max_failures:
                    raise TrainingFailedError(
                        f"training failed after {attempt} attempt(s): {exc}") from exc

    @staticmethod
    def _publish_resume_point(storage: str, attempt: int, initial=None) -> None:
        """What ``train.get_checkpoint()`` answers inside the workers: the checkpoint given to the trainer
        on the first attempt, the newest checkpoint the failed attempt reported afterwards (so a training
        loop that reports checkpoints resumes mid-run instead of starting over)."""
        marker = os.path.join(storage, ".resume_from")
        path = None
        if attempt > 0:
            found = [os.path.join(storage, d) for d in os.listdir(storage) if d.startswith("checkpoint_")]
            if found:
                path = max(found, key=os.path.getmtime)
        if path is None and initial is not None:
            path = initial.path
        if path is None:
            if attempt == 0 and os.path.exists(marker):
                os.remove(marker)
            return
        with open(marker, "w", encoding="utf-8") as fh:
            fh.write(path)

    def _result(self, reports: List[Dict[str, Any]], storage: str) -> Result:
        history = [r["metrics"] for r in reports if r["metrics"]]
        ckpts = [r["checkpoint"] for r in reports if r["checkpoint"]]
        merged: Dict[str, Any] = {}
        for m in history:
            merged.update(m)
        return Result(merged, Checkpoint(ckpts[-1]) if ckpts else None, history, storage)

    def _fit_spmd(self, run_name: str, storage: str) -> List[Dict[str, Any]]:
        import torch
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        local_rank = int(os.environ.get("LOCAL_RANK", rank))
        use_gpu = self._scaling.wants_gpu() if self._scaling.use_gpu is not None or \
            self._scaling.resources_per_worker else torch.cuda.is_available()
        shards = self._shards_for(world)[rank] if not getattr(self, "_local_shards", False) \
            else dict(self._datasets)
        backend = self._torch.resolve_backend(use_gpu)
        return run_worker(self._fn, self._config, rank, world, local_rank,
                          os.environ.get("MASTER_ADDR", "127.0.0.1"),
                          int(os.environ.get("MASTER_PORT", "29500")), backend, use_gpu, shards,
                          run_name, storage)

    def _fit_actors(self, run_name: str, storage: str) -> List[Dict[str, Any]]:
        n = self._scaling.num_workers
        use_gpu = self._scaling.wants_gpu()
        res = self._scaling.worker_resources()
        backend = self._torch.resolve_backend(use_gpu)
        if not runtime.is_initialized():
            runtime.init()
        pg = runtime.placement_group([dict(res) for _ in range(n)],
                                     strategy=self._scaling.placement_strategy)
        if not pg.wait(60):
            runtime.remove_placement_group(pg)
            raise TrainingFailedError(
                f"could not reserve {n} x {res} for training workers within 60 s "
                f"(available: {runtime.available_resources()})")
        workers = []
        try:
            custom = {k: v for k, v in res.items() if k not in ("CPU", "GPU", "memory")}
            for rank in range(n):
                w = runtime.remote(TrainWorker).options(
                    num_cpus=res.get("CPU", 1), num_gpus=res.get("GPU", 0) or None,
                    memory=res.get("memory"), resources=custom, placement_group=pg,
                    placement_group_bundle_index=rank, max_concurrency=2).remote(rank, n)
                workers.append(w)
            addr = "127.0.0.1"
            port = find_free_port()
            shards = self._shards_for(n)
            refs = [w.run.remote(self._fn, self._config, addr, port, backend, use_gpu, shards[r],
                                 run_name, storage) for r, w in enumerate(workers)]
            # Take the results in COMPLETION order: if any rank fails while the others sit in a
            # collective (a peer-flag spin, an NCCL/gloo wait), the failure must surface at once so the
            # `finally` below kills the survivors and FailureConfig can restart the group.  Waiting on
            # rank 0 first would block on a rank that can never finish.
            done: Dict[int, Any] = {}
            index = {ref._id: r for r, ref in enumerate(refs)}
            pending = list(refs)
            while pending:
                ready, pending = runtime.wait(pending, num_returns=1, timeout=None)
                for ref in ready:
                    try:
                        done[index[ref._id]] = runtime.get(ref)
                    except (runtime.TaskError, runtime.ActorDiedError) as exc:
                        # a rank that died without raising (killed, out of memory, a native crash) leaves no
                        # traceback: the tail of its log file is the evidence
                        tail = ""
                        if isinstance(exc, runtime.ActorDiedError):
                            tail = _worker_log_tail(workers[index[ref._id]])
                        raise TrainingFailedError(f"rank {index[ref._id]} failed: {exc}{tail}") from exc
            all_reports = [done[r] for r in range(n)]
            self._all_reports = all_reports                 # every rank's reports (legacy Trainer.run)
            return all_reports[0]
        finally:
            for w in workers:
                try:
                    runtime.kill(w, no_restart=True)
                except Exception:  # noqa: BLE001
                    pass
            runtime.remove_placement_group(pg)
