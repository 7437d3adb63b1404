"""XGBoost / TF estimator wrappers and the resource probe."""
import importlib.util

import numpy as np
import pytest

from raydp_b200 import runtime as rt
from raydp_b200.sql.functions import col, rand


def test_xgboost_estimator_fit_on_spark(spark_on_ray_small):
    """num_workers=2: two boosting workers, histograms all-reduced, a checkpoint per round (newest kept)."""
    import os
    from raydp_b200.xgboost import XGBoostEstimator
    spark = spark_on_ray_small
    df = spark.range(0, 4000).withColumn("x", rand(1) * 10).withColumn("y", rand(2) * 10) \
        .withColumn("z", col("x") * 3 + col("y") * 4 + 5).drop("id")
    train_df, test_df = df.randomSplit([0.8, 0.2], seed=0)
    est = XGBoostEstimator({"objective": "reg:squarederror", "eta": 0.3, "max_depth": 4}, "z",
                           num_workers=2, num_boost_round=30)
    results = est.fit_on_spark(train_df, test_df)
    assert results["features"] == ["x", "y"] and results["backend"] == "raydp_b200-hist-gbt"
    assert results["num_workers"] == 2 and results["rounds"] == 30
    rmse = results["history"]["valid"]["rmse"]
    assert len(rmse) == 30 and rmse[-1] < 1.5 and rmse[-1] < 0.2 * rmse[0]
    model = est.get_model()
    pred = model.predict(np.array([[1.0, 1.0], [5.0, 5.0]], dtype=np.float32))
    assert abs(pred[0] - 12.0) < 3.0 and abs(pred[1] - 40.0) < 3.0
    assert model.num_boosted_rounds() == 30 and len(model.get_dump()) == 30
    kept = [d for d in os.listdir(os.path.dirname(results.checkpoint.path)) if d.startswith("checkpoint_")]
    assert len(kept) == 1                                    # CheckpointConfig(num_to_keep=1)


def _gbt_data(n=3000, seed=0):
    rng = np.random.default_rng(seed)
    x = rng.standard_normal((n, 5)).astype(np.float32)
    x[rng.random((n, 5)) < 0.03] = np.nan                   # missing values take the learned default direction
    y = 2 * np.nan_to_num(x[:, 0]) - np.nan_to_num(x[:, 1]) ** 2 + (np.nan_to_num(x[:, 2]) > 0.3) * 3.0
    return x, y + 0.05 * rng.standard_normal(n)


def test_hist_gbt_single_process_learns_and_round_trips(tmp_path):
    from raydp_b200.xgboost import hist_gbt
    x, y = _gbt_data()
    model, hist = hist_gbt.train({"objective": "reg:squarederror", "eta": 0.3, "max_depth": 5, "lambda": 1.0},
                                 x[:2400], y[:2400], 40, evals=[(x[2400:], y[2400:], "valid")])
    assert hist["train"]["rmse"][-1] < 0.35 and hist["valid"]["rmse"][-1] < 0.6
    assert hist["train"]["rmse"] == sorted(hist["train"]["rmse"], reverse=True)       # squared error never goes up
    model.save_model(str(tmp_path / "m.json"))
    again = hist_gbt.HistGBTModel.load_model(str(tmp_path / "m.json"))
    assert np.allclose(again.predict(x[2400:]), model.predict(x[2400:]))
    # classification
    yb = (y > np.median(y)).astype(np.float64)
    clf, h = hist_gbt.train({"objective": "binary:logistic", "max_depth": 4, "eval_metric": ["logloss", "error"]},
                            x, yb, 25)
    assert h["train"]["error"][-1] < 0.08 and h["train"]["logloss"][-1] < 0.25
    p = clf.predict(x)
    assert p.min() >= 0 and p.max() <= 1
    with pytest.raises(ValueError, match="not supported"):
        hist_gbt.train({"booster": "gblinear"}, x, y, 1)


def test_hist_gbt_two_workers_grow_the_single_process_trees(ray_cluster, tmp_path):
    """Data parallel boosting is exact: with the same global bin edges, all-reduced histograms give the
    trees a single process grows on the whole data; a worker failure mid-way resumes from the last round."""
    import os
    import pyarrow as pa
    from raydp_b200 import data as rdata
    from raydp_b200.xgboost import XGBoostEstimator, hist_gbt
    x, y = _gbt_data(2000, seed=4)
    params = {"objective": "reg:squarederror", "eta": 0.2, "max_depth": 4, "max_bin": 64, "base_score": 0.5}
    cols = {f"f{i}": x[:, i] for i in range(5)}
    cols["y"] = y
    ds = rdata.from_arrow([pa.table({k: v[:1000] for k, v in cols.items()}),
                           pa.table({k: v[1000:] for k, v in cols.items()})])
    est = XGBoostEstimator(params, "y", num_workers=2, shuffle=False, num_boost_round=12)
    res = est.fit(ds)
    dist_model = est.get_model()
    single, hist = hist_gbt.train(params, x, y, 12)
    # same cuts are not guaranteed (the distributed sketch merges per-shard quantiles), so compare quality
    # and structure sizes rather than bits
    assert abs(res["history"]["train"]["rmse"][-1] - hist["train"]["rmse"][-1]) < 0.05
    assert np.corrcoef(dist_model.predict(x), single.predict(x))[0, 1] > 0.995
    # failure injection: rank 1 dies once after round 5; training resumes at round 6, not at 0
    flag = str(tmp_path / "died")
    real_train = hist_gbt.train

    def flaky_train(p, xx, yy, rounds, **kw):
        cb = kw["callback"]

        def wrapped(rnd, model, metrics):
            cb(rnd, model, metrics)
            from raydp_b200.train import get_context
            if rnd == 5 and get_context().get_world_rank() == 1 and not os.path.exists(flag):
                open(flag, "w").close()
                os._exit(1)
        kw["callback"] = wrapped
        return real_train(p, xx, yy, rounds, **kw)
    est2 = XGBoostEstimator(dict(params, _test_hook=None), "y", num_workers=2, shuffle=False, num_boost_round=12)
    est2._xgboost_params.pop("_test_hook")
    est2._train_fn = flaky_train
    res2 = est2.fit(ds, max_retries=2)
    assert os.path.exists(flag) and res2["rounds"] == 12
    iterations = [m["training_iteration"] for m in res2.metrics_history]
    assert iterations[0] == 7 and iterations[-1] == 12           # the second attempt started after round 6
    assert abs(res2["history"]["train"]["rmse"][-1] - res["history"]["train"]["rmse"][-1]) < 1e-6


def test_tf_estimator_is_import_guarded():
    from raydp_b200.tf import TFEstimator
    if importlib.util.find_spec("tensorflow") is None:
        with pytest.raises(ImportError, match="tensorflow"):
            TFEstimator(model=None, optimizer="adam", loss="mse", feature_columns=["x"], label_columns="y")


def test_cluster_resources(ray_cluster):
    from raydp_b200.ray_cluster_resources import ClusterResources
    assert ClusterResources.total_alive_nodes() == 1
    assert len(ClusterResources.satisfy({"CPU": 2})) == 1
    assert ClusterResources.satisfy({"CPU": 10_000}) == []
    assert rt.cluster_resources()["CPU"] == 6


def test_device_side_quantiles_and_binning_match_the_host_code():
    """The GPU booster computes its quantile proposals and bin ids with torch ops on the device; the same helpers
    on CPU tensors must reproduce numpy's quantiles and the host binning bit for bit."""
    import numpy as np
    import torch
    from raydp_b200.xgboost import hist_gbt
    rng = np.random.default_rng(0)
    x = rng.standard_normal((5000, 4)).astype(np.float32)
    x[rng.random(x.shape) < 0.05] = np.nan
    qs = np.linspace(0, 1, 2 * 16 + 1)[1:-1]
    got = hist_gbt._local_quantiles_gpu(torch.from_numpy(x), qs)
    for f in range(4):
        col = x[:, f]
        col = col[~np.isnan(col)]
        assert got[f][1] == len(col) and np.allclose(got[f][0], np.quantile(col.astype(np.float64), qs), atol=1e-6)
    cuts = hist_gbt._global_cuts(x, 16, hist_gbt._Comm())
    binned, n_bins = hist_gbt._bin(x, cuts)
    rows = hist_gbt._GpuRows(None, max(len(c) for c in cuts) + 2, torch.device("cpu"), x=torch.from_numpy(x), cuts=cuts)
    assert rows.n_bins == n_bins and np.array_equal(rows.binned_t.numpy().T.astype(np.uint16), binned)


def test_xgboost_ray_surface_trains_on_two_actors_and_predicts(ray_cluster):
    """``RayDMatrix`` from numpy / pandas / a Dataset, ``train`` with ``RayParams(num_actors=2)`` filling
    ``evals_result`` under the caller's names, ``predict``."""
    import numpy as np
    import pandas as pd
    from raydp_b200.xgboost import ray_api
    rng = np.random.default_rng(0)
    x = rng.standard_normal((4000, 5)).astype(np.float32)
    y = (x[:, 0] * 2 - x[:, 1] + 0.1 * rng.standard_normal(4000)).astype(np.float32)
    dtrain = ray_api.RayDMatrix(x[:3000], label=y[:3000])
    frame = pd.DataFrame({f"f{i}": x[3000:, i] for i in range(5)})
    frame["target"] = y[3000:]
    dtest = ray_api.RayDMatrix(frame.rename(columns={"target": "__label__"}), label="__label__")
    evals_result, extra = {}, {}
    bst = ray_api.train({"objective": "reg:squarederror", "max_depth": 4, "eval_metric": ["rmse", "mae"]}, dtrain,
                        num_boost_round=8, evals=[(dtrain, "train"), (dtest, "holdout")], evals_result=evals_result,
                        additional_results=extra, ray_params=ray_api.RayParams(num_actors=2, cpus_per_actor=1))
    assert set(evals_result) == {"train", "holdout"} and len(evals_result["holdout"]["rmse"]) == 8
    assert evals_result["holdout"]["rmse"][-1] < 0.5 * evals_result["holdout"]["rmse"][0]
    assert extra["num_actors"] == 2 and extra["total_n"] == 3000
    pred = ray_api.predict(bst, dtest)
    assert pred.shape == (1000,) and np.sqrt(np.mean((pred - y[3000:]) ** 2)) < 1.1 * evals_result["holdout"]["rmse"][-1] + 1e-6
