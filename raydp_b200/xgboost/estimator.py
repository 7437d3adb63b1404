"""``XGBoostEstimator``: gradient-boosted trees from Spark data, trained data-parallel.

Parity: reference ``python/raydp/xgboost/estimator.py:29-119`` -- ctor ``(xgboost_params,
label_column, num_workers=1, resources_per_worker=None, shuffle=True)``,
``fit(train_ds, evaluate_ds=None, max_retries=3)``, ``fit_on_spark(train_df, evaluate_df=None,
max_retries=3, fs_directory=None, compression=None, stop_spark_after_conversion=False)`` which
converts with ``parallelism=num_workers`` (``:105-112``), ``get_model()``.  The reference runs
``XGBoostTrainer`` over ``num_workers`` workers with ``CheckpointConfig(checkpoint_frequency=1,
num_to_keep=1)`` and ``FailureConfig(max_failures=max_retries)`` (``:58-81``).

Here the boosting runs on this framework's own worker group (``TorchTrainer`` actors, gloo):
:mod:`raydp_b200.xgboost.hist_gbt` grows every tree from histograms that are all-reduced across the
``num_workers`` row shards -- the algorithm of XGBoost's distributed ``hist`` method -- rank 0 writes a
checkpoint after EVERY boosting round (only the newest is kept) and a worker group restarted by
``FailureConfig`` resumes after the last checkpointed round.  When the ``xgboost`` package is importable
and ``num_workers == 1`` the package itself trains (same parameters, ``model.json`` in xgboost's own
format); ``backend`` in the result says which one ran.
"""
from __future__ import annotations

import logging
import os
import tempfile
from typing import Any, Dict, List, Optional

import numpy as np

from raydp_b200 import runtime
from raydp_b200.estimator import EstimatorInterface
from raydp_b200.spark.interfaces import SparkEstimatorInterface
from raydp_b200.train.checkpoint import Checkpoint

logger = logging.getLogger(__name__)


def _to_xy(table, label_column: str):
    cols = [c for c in table.column_names if c != label_column]
    x = np.column_stack([table.column(c).to_numpy(zero_copy_only=False) for c in cols]).astype(np.float32)
    y = table.column(label_column).to_numpy(zero_copy_only=False)
    return x, y, cols


def _have_xgboost() -> bool:
    import importlib.util
    return importlib.util.find_spec("xgboost") is not None


def _train_task(params: Dict[str, Any], label_column: str, num_boost_round: int, train_table, eval_table,
                out_dir: str) -> Dict[str, Any]:
    """Single-worker training with the xgboost package itself."""
    import xgboost as xgb  # type: ignore
    x, y, cols = _to_xy(train_table, label_column)
    os.makedirs(out_dir, exist_ok=True)
    dtrain = xgb.DMatrix(x, label=y, feature_names=cols)
    evals = [(dtrain, "train")]
    if eval_table is not None:
        xe, ye, _ = _to_xy(eval_table, label_column)
        evals.append((xgb.DMatrix(xe, label=ye, feature_names=cols), "valid"))
    history: Dict[str, Any] = {}
    booster = xgb.train(params, dtrain, num_boost_round=num_boost_round, evals=evals,
                        evals_result=history, verbose_eval=False)
    booster.save_model(os.path.join(out_dir, "model.json"))
    return {"features": cols, "backend": "xgboost", "history": history}


def _boost_loop(config: Dict[str, Any]) -> None:
    """Body of one boosting worker (rank r of num_workers): boost on the local row shard, histograms
    all-reduced across the group, one checkpoint per round from rank 0, resume after a restart."""
    from raydp_b200.train import get_checkpoint, get_context, get_dataset_shard, report
    from raydp_b200.xgboost import hist_gbt
    label = config["label_column"]
    x, y, cols = _to_xy(get_dataset_shard("train").to_arrow(), label)
    evals = []
    if config["evaluate"]:
        xe, ye, _ = _to_xy(get_dataset_shard("evaluate").to_arrow(), label)
        evals.append((xe, ye, "valid"))
    rank = get_context().get_world_rank()
    resume = None
    ckpt = get_checkpoint()
    if ckpt is not None and os.path.exists(os.path.join(ckpt.path, MODEL_FILE)):
        resume = hist_gbt.HistGBTModel.load_model(os.path.join(ckpt.path, MODEL_FILE))

    def after_round(rnd: int, model, metrics: Dict[str, float]) -> None:
        checkpoint = None
        tmp = None
        if rank == 0:
            tmp = tempfile.mkdtemp(prefix="raydp_b200_gbt_")
            model.save_model(os.path.join(tmp, MODEL_FILE))
            checkpoint = Checkpoint.from_directory(tmp)
        report(dict(metrics, training_iteration=rnd + 1, features=cols,
                    tree_method=model.attributes.get("tree_method", "hist")), checkpoint=checkpoint)
        if tmp is not None:
            import shutil
            shutil.rmtree(tmp, ignore_errors=True)

    train_fn = config.get("train_fn") or hist_gbt.train
    train_fn(config["params"], x, y, config["num_boost_round"], evals=evals, feature_names=cols,
             resume=resume, callback=after_round)


MODEL_FILE = "model.gbt.json"


class BoostResult(dict):
    """What ``fit`` returns: a dict (``features``, ``backend``, ``history`` ...) that also carries the
    trainer result's ``checkpoint`` / ``metrics`` / ``metrics_history`` attributes."""
    checkpoint = None
    metrics: Dict[str, Any] = {}
    metrics_history: List[Dict[str, Any]] = []


class XGBoostEstimator(EstimatorInterface, SparkEstimatorInterface):
    def __init__(self, xgboost_params: Dict[str, Any], label_column: str, dmatrix_params: Optional[Dict] = None,
                 num_workers: int = 1, resources_per_worker: Optional[Dict[str, float]] = None,
                 shuffle: bool = True, num_boost_round: int = 10):
        self._xgboost_params = dict(xgboost_params or {})
        self._label_column = label_column
        self._dmatrix_params = dmatrix_params
        self._num_workers = num_workers
        self._resources_per_worker = resources_per_worker
        self._shuffle = shuffle
        self._num_boost_round = int(self._xgboost_params.pop("num_boost_round", num_boost_round))
        self._results: Optional[Dict[str, Any]] = None
        self._checkpoint: Optional[Checkpoint] = None

    def fit(self, train_ds, evaluate_ds=None, max_retries: int = 3):
        if self._shuffle:
            train_ds = train_ds.random_shuffle()
            if evaluate_ds is not None:
                evaluate_ds = evaluate_ds.random_shuffle()
        if self._num_workers <= 1 and _have_xgboost():
            return self._fit_with_xgboost_package(train_ds, evaluate_ds, max_retries)
        from raydp_b200.train import (CheckpointConfig, FailureConfig, RunConfig, ScalingConfig, TorchConfig,
                                      TorchTrainer)
        res = dict(self._resources_per_worker or {})
        # a worker that was given a GPU builds its histograms with the CUDA kernels (csrc/kernels/gbt.cu); the
        # per-level histogram all-reduce between workers stays on the host (gloo): it is a few hundred KB
        use_gpu = float(res.get("GPU", 0) or 0) > 0
        datasets = {"train": train_ds}
        if evaluate_ds is not None:
            datasets["evaluate"] = evaluate_ds
        config = {"params": self._xgboost_params, "label_column": self._label_column,
                  "num_boost_round": self._num_boost_round, "evaluate": evaluate_ds is not None,
                  "train_fn": getattr(self, "_train_fn", None)}
        trainer = TorchTrainer(_boost_loop, train_loop_config=config,
                               scaling_config=ScalingConfig(num_workers=max(1, self._num_workers), use_gpu=use_gpu,
                                                            resources_per_worker=res or None),
                               run_config=RunConfig(checkpoint_config=CheckpointConfig(checkpoint_frequency=1,
                                                                                       num_to_keep=1),
                                                    failure_config=FailureConfig(max_failures=max_retries)),
                               torch_config=TorchConfig(backend="gloo"), datasets=datasets)
        result = trainer.fit()
        self._checkpoint = result.checkpoint
        history: Dict[str, Dict[str, List[float]]] = {}
        for m in result.metrics_history:
            for key, value in m.items():
                if "-" in key and isinstance(value, float):
                    name, metric = key.split("-", 1)
                    history.setdefault(name, {}).setdefault(metric, []).append(value)
        self._results = BoostResult(features=result.metrics.get("features"), backend="raydp_b200-hist-gbt",
                                    history=history, num_workers=max(1, self._num_workers),
                                    rounds=result.metrics.get("training_iteration"),
                                    tree_method=result.metrics.get("tree_method", "hist"))
        self._results.checkpoint, self._results.metrics = result.checkpoint, result.metrics
        self._results.metrics_history = result.metrics_history
        return self._results

    def _fit_with_xgboost_package(self, train_ds, evaluate_ds, max_retries: int):
        out_dir = tempfile.mkdtemp(prefix="raydp_b200_xgb_")
        cpus = float((self._resources_per_worker or {}).get("CPU", 1))
        task = runtime.remote(_train_task).options(num_cpus=cpus, max_retries=max_retries)
        train_ref = runtime.put(train_ds.to_arrow())
        eval_ref = runtime.put(evaluate_ds.to_arrow()) if evaluate_ds is not None else None
        out = runtime.get(task.remote(self._xgboost_params, self._label_column, self._num_boost_round,
                                      train_ref, eval_ref, out_dir))
        runtime.free([r for r in (train_ref, eval_ref) if r is not None])
        self._checkpoint = Checkpoint.from_directory(out_dir)
        self._results = BoostResult(**out)
        self._results.checkpoint, self._results.metrics, self._results.metrics_history = self._checkpoint, out, [out]
        return self._results

    def fit_on_spark(self, train_df, evaluate_df=None, fs_directory: Optional[str] = None,
                     compression: Optional[str] = None, max_retries: int = 3,
                     stop_spark_after_conversion: bool = False):
        from raydp_b200.context import stop_spark
        from raydp_b200.spark.dataset import read_spark_parquet, spark_dataframe_to_ray_dataset
        train_df = self._check_and_convert(train_df)
        evaluate_ds = None
        if evaluate_df is not None:
            evaluate_df = self._check_and_convert(evaluate_df)
        if fs_directory is not None:
            app_id = train_df.sql_ctx.sparkSession.sparkContext.applicationId
            path = fs_directory.rstrip("/") + f"/{app_id}"
            train_df.write.parquet(path + "/train", compression=compression)
            train_ds = read_spark_parquet(path + "/train")
            if evaluate_df is not None:
                evaluate_df.write.parquet(path + "/test", compression=compression)
                evaluate_ds = read_spark_parquet(path + "/test")
        else:
            train_ds = spark_dataframe_to_ray_dataset(train_df, parallelism=self._num_workers)
            if evaluate_df is not None:
                evaluate_ds = spark_dataframe_to_ray_dataset(evaluate_df, parallelism=self._num_workers)
        if stop_spark_after_conversion:
            stop_spark(cleanup_data=False)
        return self.fit(train_ds, evaluate_ds, max_retries)

    def get_model(self):
        assert self._checkpoint is not None, "Must call fit first"
        path = self._checkpoint.path
        if os.path.exists(os.path.join(path, MODEL_FILE)):
            from raydp_b200.xgboost.hist_gbt import HistGBTModel
            return HistGBTModel.load_model(os.path.join(path, MODEL_FILE))
        import xgboost as xgb  # type: ignore
        booster = xgb.Booster()
        booster.load_model(os.path.join(path, "model.json"))
        return booster

    @property
    def results(self):
        return self._results
