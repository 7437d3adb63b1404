"""Gradient-boosted trees on the NYC-taxi features with ``XGBoostEstimator``.

Counterpart of the reference's ``examples/xgboost_ray_nyctaxi.py`` (Spark preprocessing, 90/10
split, ``tree_method=hist`` regression).  Two boosting workers grow every tree together (per-level histogram
all-reduce); with ``--gpu`` each worker gets a GPU and builds its histograms with the CUDA kernels
(``csrc/kernels/gbt.cu``).  The ``xgboost`` package itself is only used for ``num_workers=1`` when it is installed.

    python examples/xgboost_nyctaxi.py [--gpu]
"""
import sys
import _bootstrap  # noqa: F401
import numpy as np

import raydp
from raydp import runtime as ray
from raydp.utils import random_split
from raydp.xgboost import XGBoostEstimator

from data_process import LABEL, csv_path, nyc_taxi_preprocess

if __name__ == "__main__":
    gpu = "--gpu" in sys.argv
    ray.init(address="local", num_cpus=4, num_gpus=2 if gpu else 0)
    spark = raydp.init_spark("NYC Taxi Fare Prediction with RayDP", num_executors=1, executor_cores=1,
                             executor_memory="500M")
    frame = spark.read.format("csv").option("header", "true").option("inferSchema", "true").load(csv_path())
    spark.conf.set("spark.sql.session.timeZone", "UTC")
    frame = nyc_taxi_preprocess(frame)
    train_df, test_df = random_split(frame, [0.9, 0.1], 0)

    params = {"tree_method": "hist", "objective": "reg:squarederror", "eval_metric": ["rmse", "mae"],
              "max_depth": 6, "eta": 0.2, "num_boost_round": 40}
    resources = {"CPU": 1, "GPU": 1} if gpu else {"CPU": 1}
    estimator = XGBoostEstimator(params, LABEL, num_workers=2, resources_per_worker=resources)
    estimator.fit_on_spark(train_df, test_df)
    booster = estimator.get_model()

    holdout = test_df.toPandas()
    x = holdout.drop(columns=[LABEL]).to_numpy(dtype=np.float32)
    if hasattr(booster, "inplace_predict"):          # xgboost.Booster
        pred = booster.inplace_predict(x)
    else:                                            # the built-in booster (HistGBTModel)
        pred = booster.predict(x)
    print("model:", type(booster).__name__, "hold-out MAE:", float(np.abs(pred - holdout[LABEL].to_numpy()).mean()))
    raydp.stop_spark()
    ray.shutdown()
