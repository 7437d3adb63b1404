"""The first-generation Ray Train API: ``Trainer(backend="torch", num_workers=n).start() / run() / shutdown()``.

The reference's ``examples/raytrain_nyctaxi.py:5,167-177`` and its tutorial notebooks are written
against it (``Trainer``, ``TrainingCallback``, ``train.report(**metrics)``, ``to_tune_trainable``).  It is a
thin front over :class:`raydp_b200.train.trainer.TorchTrainer`: ``run`` launches the worker group for one
call and returns what ``train_func`` returned on every worker.
"""
from __future__ import annotations

from typing import Any, Callable, Dict, List, Optional

from raydp_b200.train import session as _session
from raydp_b200.train.config import RunConfig, ScalingConfig, TorchConfig
from raydp_b200.train.trainer import TorchTrainer

_RETURN_KEY = "__train_func_return__"


class TrainingCallback:
    """Hooks around a ``Trainer.run``: ``handle_result`` gets one list (a metrics dict per worker) for every
    ``train.report`` round."""

    def start_training(self, logdir: str = "", config: Optional[dict] = None, **info) -> None:
        pass

    def handle_result(self, results: List[Dict], **info) -> None:
        pass

    def process_results(self, results: List[Dict]) -> List[Dict]:
        return results

    def finish_training(self, error: bool = False, **info) -> None:
        pass


class Trainer:
    def __init__(self, backend: str = "torch", num_workers: int = 1, use_gpu: bool = False,
                 resources_per_worker: Optional[Dict[str, float]] = None, logdir: Optional[str] = None,
                 max_retries: int = 0):
        if str(backend).lower() not in ("torch", "pytorch"):
            raise ValueError(f"backend {backend!r} is not available: this runtime trains torch models")
        self._scaling = ScalingConfig(num_workers=num_workers, use_gpu=use_gpu,
                                      resources_per_worker=resources_per_worker)
        self._logdir = logdir
        self._started = False
        self.latest_checkpoint = None

    def start(self, initialization_hook: Optional[Callable] = None) -> None:
        self._started = True

    def shutdown(self) -> None:
        self._started = False

    def run(self, train_func: Callable, config: Optional[dict] = None, callbacks: Optional[List[TrainingCallback]] = None,
            dataset=None, checkpoint=None, checkpoint_strategy=None) -> List[Any]:
        if not self._started:
            raise RuntimeError("Trainer.start() must be called before run()")
        callbacks = list(callbacks or [])
        datasets = dataset if isinstance(dataset, dict) else ({"train": dataset} if dataset is not None else None)

        def loop(cfg=None, _fn=train_func, _has_config=config is not None):
            out = _fn(cfg) if _has_config else _fn()
            _session.get_session().reports.append({"metrics": {_RETURN_KEY: out}, "checkpoint": None})

        trainer = TorchTrainer(loop, train_loop_config=config if config is not None else None,
                               scaling_config=self._scaling,
                               run_config=RunConfig(storage_path=self._logdir), torch_config=TorchConfig(),
                               datasets=datasets)
        for cb in callbacks:
            cb.start_training(logdir=self._logdir or "", config=config or {})
        try:
            result = trainer.fit()
        except BaseException:
            for cb in callbacks:
                cb.finish_training(error=True)
            raise
        per_worker = getattr(trainer, "_all_reports", None) or [[{"metrics": m, "checkpoint": None}
                                                                 for m in result.metrics_history]]
        rounds = max((len(r) for r in per_worker), default=0)
        for i in range(rounds):
            batch = [r[i]["metrics"] for r in per_worker if i < len(r) and _RETURN_KEY not in r[i]["metrics"]]
            if batch:
                for cb in callbacks:
                    cb.handle_result(cb.process_results(batch))
        for cb in callbacks:
            cb.finish_training(error=False)
        self.latest_checkpoint = result.checkpoint
        returns = []
        for reports in per_worker:
            last = [r["metrics"][_RETURN_KEY] for r in reports if _RETURN_KEY in r["metrics"]]
            returns.append(last[-1] if last else None)
        return returns

    def to_tune_trainable(self, train_func: Callable, dataset=None) -> Callable[[dict], None]:
        """A function trainable for ``tune.run``: one ``Trainer.run`` per trial; rank 0's reported metrics
        are forwarded to the tuner."""
        scaling = self._scaling

        def trainable(config: dict) -> None:
            from raydp_b200 import tune
            inner = Trainer(num_workers=scaling.num_workers, use_gpu=bool(scaling.use_gpu),
                            resources_per_worker=scaling.resources_per_worker)
            inner.start()
            seen: List[Dict] = []

            class _Forward(TrainingCallback):
                def handle_result(self, results, **info):
                    seen.append(results[0])
            inner.run(train_func, config=config, callbacks=[_Forward()], dataset=dataset)
            inner.shutdown()
            for metrics in seen:
                tune.report(**metrics)
        return trainable
