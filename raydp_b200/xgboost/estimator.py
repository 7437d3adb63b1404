"""``XGBoostEstimator``: gradient-boosted trees from Spark data.

Parity: reference ``python/raydp/xgboost/estimator.py:29-119`` -- ctor ``(xgboost_params,
label_column, dmatrix_params=None, num_workers=1, resources_per_worker=None, shuffle=True)``,
``fit(train_ds, evaluate_ds=None, max_retries=3)``, ``fit_on_spark(train_df, evaluate_df=None,
fs_directory=None, compression=None, max_retries=3, stop_spark_after_conversion=False)`` which
converts with ``parallelism=num_workers`` (``:105-112``), checkpoint kept per iteration
(``CheckpointConfig(checkpoint_frequency=1, num_to_keep=1)`` ``:60-68``), ``get_model()``.

Backends: the ``xgboost`` package when importable (same params pass straight through); otherwise
scikit-learn's histogram gradient boosting with the common xgboost parameter names mapped
(``objective``, ``eta``/``learning_rate``, ``max_depth``, ``lambda``, ``num_boost_round``), so the
estimator stays usable in images without xgboost.  Training runs in one runtime task; the
reference's rabit all-reduce of histograms has no counterpart here (``num_workers`` only sets the
conversion parallelism).
"""
from __future__ import annotations

import logging
import os
import pickle
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


def _train_task(params: Dict[str, Any], label_column: str, num_boost_round: int, train_table, eval_table,
                out_dir: str) -> Dict[str, Any]:
    x, y, cols = _to_xy(train_table, label_column)
    result: Dict[str, Any] = {"features": cols}
    os.makedirs(out_dir, exist_ok=True)
    try:
        import xgboost as xgb  # type: ignore
        dtrain = xgb.DMatrix(x, label=y, feature_names=cols)
        evals = [(dtrain, "train")]
        if eval_table is not None:
            xe, ye, _ = _to_xy(eval_table, label_column)
            evals.append((xgb.DMatrix(xe, label=ye, feature_names=cols), "valid"))
        history: Dict[str, Any] = {}
        booster = xgb.train(params, dtrain, num_boost_round=num_boost_round, evals=evals,
                            evals_result=history, verbose_eval=False)
        booster.save_model(os.path.join(out_dir, "model.json"))
        result.update(backend="xgboost", history=history)
        return result
    except ImportError:
        pass
    from sklearn.ensemble import HistGradientBoostingClassifier, HistGradientBoostingRegressor
    objective = str(params.get("objective", "reg:squarederror"))
    common = dict(learning_rate=float(params.get("eta", params.get("learning_rate", 0.3))),
                  max_depth=int(params["max_depth"]) if params.get("max_depth") else None,
                  max_iter=int(num_boost_round), l2_regularization=float(params.get("lambda", 0.0)),
                  random_state=int(params.get("seed", 0)))
    if objective.startswith(("binary:", "multi:")):
        model = HistGradientBoostingClassifier(**common)
    else:
        loss = "absolute_error" if objective == "reg:absoluteerror" else "squared_error"
        model = HistGradientBoostingRegressor(loss=loss, **common)
    model.fit(x, y)
    metrics = {"train-score": float(model.score(x, y))}
    if eval_table is not None:
        xe, ye, _ = _to_xy(eval_table, label_column)
        metrics["valid-score"] = float(model.score(xe, ye))
    with open(os.path.join(out_dir, "model.pkl"), "wb") as fh:
        pickle.dump(model, fh)
    result.update(backend="sklearn-hist-gbt", history=metrics)
    return result


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
        if self._num_workers > 1:
            logger.warning("XGBoostEstimator trains in one process here; num_workers=%d only sets the "
                           "conversion parallelism", self._num_workers)
        out_dir = tempfile.mkdtemp(prefix="raydp_b200_xgb_")
        cpus = float((self._resources_per_worker or {}).get("CPU", 1))
        task = runtime.remote(_train_task).options(num_cpus=cpus, max_retries=max_retries)
        train_ref = runtime.put(train_ds.to_arrow())
        eval_ref = runtime.put(evaluate_ds.to_arrow()) if evaluate_ds is not None else None
        self._results = runtime.get(task.remote(self._xgboost_params, self._label_column, self._num_boost_round,
                                                train_ref, eval_ref, out_dir))
        runtime.free([r for r in (train_ref, eval_ref) if r is not None])
        self._checkpoint = Checkpoint.from_directory(out_dir)
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
        if os.path.exists(os.path.join(path, "model.json")):
            import xgboost as xgb  # type: ignore
            booster = xgb.Booster()
            booster.load_model(os.path.join(path, "model.json"))
            return booster
        with open(os.path.join(path, "model.pkl"), "rb") as fh:
            model = pickle.load(fh)
        # code written for xgboost.Booster calls inplace_predict(ndarray); give the fallback model that name
        if not hasattr(model, "inplace_predict"):
            model.inplace_predict = model.predict
        return model

    @property
    def results(self):
        return self._results
