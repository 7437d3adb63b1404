"""Hyper-parameter search over training runs: the slice of ``ray.tune`` that RayDP programs use.

The reference's ray-train example ends with ``tune.run(trainer.to_tune_trainable(...), config={...,
"lr": tune.grid_search([...])})`` / ``analysis.get_best_config(metric, mode)``
(``examples/raytrain_nyctaxi.py:179-190``) and its estimator signature mentions ``tune`` sample domains
(``torch/estimator.py:39,93``).  Here: sample domains, grid expansion, ``run`` and ``Tuner.fit``.  A trial
is a function ``trainable(config)`` that calls ``tune.report(...)`` (or ``train.report`` /
``session.report``), or a ``TorchTrainer`` whose ``train_loop_config`` is searched.  Trials are
independent, so they run as parallel tasks on the runtime when one is up (one CPU each by default)
and in this process otherwise.
"""
from __future__ import annotations

import copy
import itertools
import math
import random
from typing import Any, Callable, Dict, List, Optional

from raydp_b200.train import session as _session


# ------------------------------------------------------------------------------------------ search space
class Domain:
    """A distribution a parameter is drawn from (``ray.tune.search.sample.Domain``)."""

    def sample(self, rng: random.Random):
        raise NotImplementedError


class _Fn(Domain):
    def __init__(self, draw: Callable[[random.Random], Any], text: str):
        self._draw, self._text = draw, text

    def sample(self, rng: random.Random):
        return self._draw(rng)

    def __repr__(self):
        return self._text


def uniform(lower: float, upper: float) -> Domain:
    return _Fn(lambda r: r.uniform(lower, upper), f"uniform({lower}, {upper})")


def quniform(lower: float, upper: float, q: float) -> Domain:
    return _Fn(lambda r: round(r.uniform(lower, upper) / q) * q, f"quniform({lower}, {upper}, {q})")


def loguniform(lower: float, upper: float, base: float = 10) -> Domain:
    lo, hi = math.log(lower, base), math.log(upper, base)
    return _Fn(lambda r: base ** r.uniform(lo, hi), f"loguniform({lower}, {upper})")


def randint(lower: int, upper: int) -> Domain:
    return _Fn(lambda r: r.randrange(lower, upper), f"randint({lower}, {upper})")


def randn(mean: float = 0.0, sd: float = 1.0) -> Domain:
    return _Fn(lambda r: r.gauss(mean, sd), f"randn({mean}, {sd})")


def choice(categories) -> Domain:
    options = list(categories)
    return _Fn(lambda r: r.choice(options), f"choice({options})")


def sample_from(func: Callable) -> Domain:
    return _Fn(lambda r: func(None), "sample_from(...)")


def grid_search(values) -> Dict[str, list]:
    return {"grid_search": list(values)}


def _expand(space: Any, rng: random.Random) -> List[Any]:
    """All grid combinations of ``space`` with every Domain sampled once per combination."""
    if isinstance(space, dict) and set(space) == {"grid_search"}:
        return [v for option in space["grid_search"] for v in _expand(option, rng)]
    if isinstance(space, dict):
        keys = list(space)
        per_key = [_expand(space[k], rng) for k in keys]
        return [dict(zip(keys, combo)) for combo in itertools.product(*per_key)]
    if isinstance(space, Domain):
        return [space]                                  # sampled later, once per trial
    return [space]


def _sample(value: Any, rng: random.Random) -> Any:
    if isinstance(value, Domain):
        return value.sample(rng)
    if isinstance(value, dict):
        return {k: _sample(v, rng) for k, v in value.items()}
    return value


def generate_configs(space: Optional[dict], num_samples: int = 1, seed: Optional[int] = None) -> List[dict]:
    rng = random.Random(seed)
    out = []
    for _ in range(max(1, num_samples)):
        for combo in _expand(space or {}, rng):
            out.append(_sample(copy.deepcopy(combo), rng))
    return out


# ------------------------------------------------------------------------------------------ reporting
def report(_metrics: Optional[Dict[str, Any]] = None, checkpoint=None, **kwargs) -> None:
    """``tune.report(loss=...)`` / ``tune.report({"loss": ...})`` from inside a trial."""
    _session.report(_metrics, checkpoint, **kwargs)


def get_trial_name() -> str:
    return _session.get_session().run_name


def with_parameters(trainable: Callable, **params) -> Callable:
    def bound(config, _t=trainable, _p=params):
        return _t(config, **_p)
    bound.__name__ = getattr(trainable, "__name__", "trainable")
    return bound


def with_resources(trainable, resources) -> Any:
    return trainable


# ------------------------------------------------------------------------------------------ trials
class Trial:
    def __init__(self, trial_id: str, config: dict):
        self.trial_id, self.config = trial_id, config
        self.metrics_history: List[Dict[str, Any]] = []
        self.last_result: Dict[str, Any] = {}
        self.checkpoint = None
        self.error: Optional[str] = None
        self.status = "PENDING"

    def metric(self, name: str, mode: str, scope: str = "last"):
        values = [m[name] for m in self.metrics_history if name in m]
        if not values:
            return None
        if scope == "all":
            return min(values) if mode == "min" else max(values)
        return values[-1]

    def __repr__(self):
        return f"Trial({self.trial_id}, {self.status}, config={self.config})"


def _run_function_trial(trainable: Callable, config: dict, trial_id: str, storage: str):
    """Body of one trial (runs in a runtime task or in the caller's process)."""
    import torch
    previous = _session._process_session
    sess = _session.TrainSession(0, 1, 0, 1, {}, torch.device("cpu"), trial_id, storage, "gloo")
    _session._set_session(sess)
    try:
        out = trainable(config)
        if isinstance(out, dict):                       # "return metrics" style trainables
            sess.report(out)
        return sess.reports, None
    except Exception as exc:  # noqa: BLE001  (a failed trial is recorded, the search goes on)
        import traceback
        return sess.reports, f"{type(exc).__name__}: {exc}\n{traceback.format_exc()}"
    finally:
        _session._set_session(previous)


def _execute(trainable, configs: List[dict], name: str, storage: Optional[str], resources: Optional[dict],
             max_concurrent: Optional[int], raise_on_failed_trial: bool) -> List[Trial]:
    import os
    import tempfile
    from raydp_b200 import runtime
    from raydp_b200.train.trainer import TorchTrainer
    base = os.path.join(storage or os.path.join(tempfile.gettempdir(), "raydp_b200_results"), name)
    os.makedirs(base, exist_ok=True)
    trials = [Trial(f"{name}_{i:05d}", cfg) for i, cfg in enumerate(configs)]

    def finish(trial: Trial, reports, error) -> None:
        trial.metrics_history = [dict(r["metrics"], config=trial.config, trial_id=trial.trial_id) for r in reports
                                 if r["metrics"]]
        trial.last_result = trial.metrics_history[-1] if trial.metrics_history else {}
        ckpts = [r["checkpoint"] for r in reports if r.get("checkpoint")]
        if ckpts:
            from raydp_b200.train.checkpoint import Checkpoint
            trial.checkpoint = Checkpoint(ckpts[-1])
        trial.error, trial.status = error, ("ERROR" if error else "TERMINATED")
        if error and raise_on_failed_trial:
            raise TuneError(f"trial {trial.trial_id} failed:\n{error}")

    if isinstance(trainable, TorchTrainer):
        for trial in trials:                            # each trial is itself a distributed run
            overrides = trial.config.get("train_loop_config", trial.config)
            cfg = dict(trainable._config or {})
            cfg.update(overrides)
            run = TorchTrainer(trainable._fn, train_loop_config=cfg, scaling_config=trainable._scaling,
                               run_config=trainable._run, torch_config=trainable._torch,
                               datasets=trainable._datasets)
            try:
                result = run.fit()
                reports = [{"metrics": m, "checkpoint": None} for m in result.metrics_history]
                if result.checkpoint is not None and reports:
                    reports[-1]["checkpoint"] = result.checkpoint.path
                finish(trial, reports, None)
            except Exception as exc:  # noqa: BLE001
                finish(trial, [], f"{type(exc).__name__}: {exc}")
        return trials
    cpus = float((resources or {}).get("cpu", (resources or {}).get("CPU", 1)))
    if runtime.is_initialized() and not _session._process_session:
        task = runtime.remote(num_cpus=cpus)(_run_function_trial)
        width = max_concurrent or len(trials)
        for start in range(0, len(trials), width):
            wave = trials[start:start + width]
            refs = [task.remote(trainable, t.config, t.trial_id, base) for t in wave]
            for trial, (reports, error) in zip(wave, runtime.get(refs)):
                finish(trial, reports, error)
    else:
        for trial in trials:
            reports, error = _run_function_trial(trainable, trial.config, trial.trial_id, base)
            finish(trial, reports, error)
    return trials


class TuneError(RuntimeError):
    pass


# ------------------------------------------------------------------------------------------ results
def _best(trials: List[Trial], metric: str, mode: str, scope: str = "last") -> Optional[Trial]:
    if mode not in ("min", "max"):
        raise ValueError("mode must be 'min' or 'max'")
    scored = [(t.metric(metric, mode, scope), t) for t in trials]
    scored = [(v, t) for v, t in scored if v is not None]
    if not scored:
        return None
    pick = min if mode == "min" else max
    return pick(scored, key=lambda vt: vt[0])[1]


class ExperimentAnalysis:
    """What ``tune.run`` returns."""

    def __init__(self, trials: List[Trial], default_metric: Optional[str], default_mode: Optional[str]):
        self.trials = trials
        self.default_metric, self.default_mode = default_metric, default_mode

    def _resolve(self, metric, mode):
        metric, mode = metric or self.default_metric, mode or self.default_mode
        if metric is None or mode is None:
            raise ValueError("pass metric= and mode= (or give them to tune.run)")
        return metric, mode

    def get_best_trial(self, metric=None, mode=None, scope: str = "last") -> Optional[Trial]:
        return _best(self.trials, *self._resolve(metric, mode), scope)

    def get_best_config(self, metric=None, mode=None, scope: str = "last") -> Optional[dict]:
        best = self.get_best_trial(metric, mode, scope)
        return None if best is None else best.config

    @property
    def best_trial(self):
        return self.get_best_trial()

    @property
    def best_config(self):
        return self.get_best_config()

    @property
    def best_result(self):
        best = self.get_best_trial()
        return None if best is None else best.last_result

    @property
    def results(self) -> Dict[str, Dict[str, Any]]:
        return {t.trial_id: t.last_result for t in self.trials}

    def dataframe(self, metric=None, mode=None):
        import pandas as pd
        rows = []
        for t in self.trials:
            row = {k: v for k, v in t.last_result.items() if k != "config"}
            row.update({f"config/{k}": v for k, v in _flatten(t.config).items()})
            rows.append(row)
        return pd.DataFrame(rows)


def _flatten(config: dict, prefix: str = "") -> Dict[str, Any]:
    out: Dict[str, Any] = {}
    for k, v in config.items():
        if isinstance(v, dict):
            out.update(_flatten(v, f"{prefix}{k}/"))
        else:
            out[f"{prefix}{k}"] = v
    return out


def run(run_or_experiment, config: Optional[dict] = None, num_samples: int = 1, metric: Optional[str] = None,
        mode: Optional[str] = None, name: Optional[str] = None, resources_per_trial: Optional[dict] = None,
        storage_path: Optional[str] = None, local_dir: Optional[str] = None, max_concurrent_trials: Optional[int] = None,
        raise_on_failed_trial: bool = True, verbose: int = 0, seed: Optional[int] = None, **_ignored) -> ExperimentAnalysis:
    import time
    name = name or f"tune_{time.strftime('%Y%m%d_%H%M%S')}"
    configs = generate_configs(config, num_samples, seed)
    trials = _execute(run_or_experiment, configs, name, storage_path or local_dir, resources_per_trial,
                      max_concurrent_trials, raise_on_failed_trial)
    return ExperimentAnalysis(trials, metric, mode)


class TuneConfig:
    def __init__(self, metric: Optional[str] = None, mode: Optional[str] = None, num_samples: int = 1,
                 max_concurrent_trials: Optional[int] = None, search_alg=None, scheduler=None, seed: Optional[int] = None,
                 **_ignored):
        self.metric, self.mode, self.num_samples = metric, mode, num_samples
        self.max_concurrent_trials, self.seed = max_concurrent_trials, seed


class ResultGrid:
    def __init__(self, trials: List[Trial], metric: Optional[str], mode: Optional[str], storage: str):
        from raydp_b200.train.trainer import Result
        self._trials, self._metric, self._mode = trials, metric, mode
        self._results = [Result({k: v for k, v in t.last_result.items()}, t.checkpoint, t.metrics_history, storage,
                                RuntimeError(t.error) if t.error else None) for t in trials]
        for r, t in zip(self._results, trials):
            r.config = t.config

    def __len__(self) -> int:
        return len(self._results)

    def __getitem__(self, i: int):
        return self._results[i]

    def __iter__(self):
        return iter(self._results)

    @property
    def errors(self) -> list:
        return [r.error for r in self._results if r.error is not None]

    @property
    def num_errors(self) -> int:
        return len(self.errors)

    def get_best_result(self, metric: Optional[str] = None, mode: Optional[str] = None, scope: str = "last"):
        metric, mode = metric or self._metric, mode or self._mode
        if metric is None or mode is None:
            raise ValueError("pass metric= and mode= (or set them in TuneConfig)")
        best = _best(self._trials, metric, mode, scope)
        if best is None:
            raise RuntimeError(f"no trial reported the metric {metric!r}")
        return self._results[self._trials.index(best)]

    def get_dataframe(self):
        return ExperimentAnalysis(self._trials, self._metric, self._mode).dataframe()


class Tuner:
    def __init__(self, trainable=None, *, param_space: Optional[dict] = None, tune_config: Optional[TuneConfig] = None,
                 run_config=None):
        self._trainable, self._space = trainable, param_space
        self._tune, self._run = tune_config or TuneConfig(), run_config

    def fit(self) -> ResultGrid:
        import time
        name = getattr(self._run, "name", None) or f"tuner_{time.strftime('%Y%m%d_%H%M%S')}"
        storage = getattr(self._run, "storage_path", None)
        configs = generate_configs(self._space, self._tune.num_samples, self._tune.seed)
        trials = _execute(self._trainable, configs, name, storage, None, self._tune.max_concurrent_trials, False)
        return ResultGrid(trials, self._tune.metric, self._tune.mode, storage or "")
