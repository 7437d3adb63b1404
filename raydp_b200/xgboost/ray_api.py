"""The ``xgboost_ray`` names the reference's example is written against (``examples/xgboost_ray_nyctaxi.py:5,58-72``:
``RayDMatrix``, ``RayParams``, ``train``), served by this framework's distributed booster.

``raydp_b200.compat`` registers this module as ``xgboost_ray`` when the real package is absent, so

    from xgboost_ray import RayDMatrix, train, RayParams
    bst = train(config, RayDMatrix(train_dataset, label="y"), evals=[(dtest, "eval")], evals_result=res,
                ray_params=RayParams(num_actors=2, cpus_per_actor=1), num_boost_round=10)

runs as written: ``num_actors`` boosting workers on the trainer's worker group, one histogram all-reduce per tree
level, GPU histogram kernels when ``gpus_per_actor`` > 0.  The returned booster is a
:class:`raydp_b200.xgboost.hist_gbt.HistGBTModel` (``predict``, ``save_model``, ``get_dump``).
"""
from __future__ import annotations

from typing import Any, Dict, List, Optional, Sequence, Tuple

import numpy as np


class RayParams:
    def __init__(self, num_actors: int = 1, cpus_per_actor: int = 1, gpus_per_actor: int = 0,
                 resources_per_actor: Optional[Dict[str, float]] = None, max_actor_restarts: int = 0,
                 max_failed_actors: int = 0, elastic_training: bool = False, **_ignored):
        if elastic_training:
            raise NotImplementedError("elastic_training is not supported: a failed worker restarts the group, which "
                                      "resumes from the last round's checkpoint (max_actor_restarts)")
        self.num_actors = int(num_actors)
        self.cpus_per_actor = cpus_per_actor
        self.gpus_per_actor = gpus_per_actor
        self.resources_per_actor = dict(resources_per_actor or {})
        self.max_actor_restarts = int(max_actor_restarts)

    def resources(self) -> Dict[str, float]:
        res = dict(self.resources_per_actor)
        if self.cpus_per_actor:
            res["CPU"] = float(self.cpus_per_actor)
        if self.gpus_per_actor and self.gpus_per_actor > 0:
            res["GPU"] = float(self.gpus_per_actor)
        return res


class RayDMatrix:
    """Training / evaluation data plus the name (or values) of the label: a :class:`raydp_b200.data.Dataset`, a
    pandas DataFrame, an Arrow table, or a numpy matrix (then ``label`` is an array)."""

    def __init__(self, data, label=None, feature_names: Optional[Sequence[str]] = None, **_ignored):
        import pyarrow as pa
        from raydp_b200 import data as rdata
        self.label = label
        if isinstance(data, rdata.Dataset):
            self.dataset = data
        elif isinstance(data, pa.Table):
            self.dataset = rdata.from_arrow(data)
        elif isinstance(data, np.ndarray):
            names = list(feature_names) if feature_names is not None else [f"f{i}" for i in range(data.shape[1])]
            cols = {n: data[:, i] for i, n in enumerate(names)}
            if label is not None and not isinstance(label, str):
                cols["__label__"] = np.asarray(label)
                self.label = "__label__"
            self.dataset = rdata.from_arrow(pa.table(cols))
        else:                                             # pandas
            frame = data
            if label is not None and not isinstance(label, str):
                frame = frame.copy()
                frame["__label__"] = np.asarray(label)
                self.label = "__label__"
            self.dataset = rdata.from_pandas(frame)
        if label is not None and not isinstance(self.label, str):
            raise TypeError("label must be a column name, or an array next to a numpy / pandas input")


def train(params: Dict[str, Any], dtrain: RayDMatrix, num_boost_round: int = 10,
          evals: Sequence[Tuple[RayDMatrix, str]] = (), evals_result: Optional[Dict[str, Dict[str, List[float]]]] = None,
          additional_results: Optional[Dict[str, Any]] = None, ray_params: Optional[RayParams] = None, **_ignored):
    """``xgboost_ray.train``: boost on ``ray_params.num_actors`` workers; fills ``evals_result`` like xgboost does
    (``{eval name: {metric: [value per round]}}``) and returns the booster."""
    from raydp_b200.xgboost.estimator import XGBoostEstimator
    rp = ray_params or RayParams()
    if dtrain.label is None:
        raise ValueError("the training RayDMatrix needs a label")
    held_out = [(dm, name) for dm, name in evals if dm is not dtrain]
    if len(held_out) > 1:
        raise NotImplementedError("one evaluation set besides the training data is supported")
    est = XGBoostEstimator(dict(params), dtrain.label, num_workers=rp.num_actors,
                           resources_per_worker=rp.resources() or None, shuffle=False, num_boost_round=num_boost_round)
    result = est.fit(dtrain.dataset, held_out[0][0].dataset if held_out else None,
                     max_retries=rp.max_actor_restarts)
    history = result.get("history") or {}
    if evals_result is not None:
        for dm, name in evals:
            src = "train" if dm is dtrain else "valid"
            if src in history:
                evals_result[name] = {m: list(v) for m, v in history[src].items()}
    if additional_results is not None:
        additional_results.update(total_n=dtrain.dataset.count(), num_actors=rp.num_actors,
                                  tree_method=result.get("tree_method"))
    return est.get_model()


def predict(model, data: RayDMatrix, ray_params: Optional[RayParams] = None, **_ignored) -> np.ndarray:
    """``xgboost_ray.predict``: scores of every row of ``data`` (computed in this process, block by block)."""
    from raydp_b200.xgboost.estimator import _to_xy
    outs = []
    for table in data.dataset.iter_tables():
        if table.num_rows == 0:
            continue
        if data.label is not None and data.label in table.column_names:
            x, _, _ = _to_xy(table, data.label)
        else:
            x = np.stack([np.asarray(table.column(n).to_numpy(zero_copy_only=False), dtype=np.float32)
                          for n in table.column_names], axis=1)
        outs.append(model.predict(x))
    return np.concatenate(outs) if outs else np.zeros(0)
