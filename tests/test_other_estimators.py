"""XGBoost / TF estimator wrappers and the resource probe."""
import importlib.util

import numpy as np
import pytest

from raydp_b200 import runtime as rt
from raydp_b200.sql.functions import col, rand


def test_xgboost_estimator_fit_on_spark(spark_on_ray_small):
    from raydp_b200.xgboost import XGBoostEstimator
    spark = spark_on_ray_small
    df = spark.range(0, 4000).withColumn("x", rand(1) * 10).withColumn("y", rand(2) * 10) \
        .withColumn("z", col("x") * 3 + col("y") * 4 + 5).drop("id")
    train_df, test_df = df.randomSplit([0.8, 0.2], seed=0)
    est = XGBoostEstimator({"objective": "reg:squarederror", "eta": 0.3, "max_depth": 4}, "z",
                           num_workers=2, num_boost_round=30)
    results = est.fit_on_spark(train_df, test_df)
    assert results["features"] == ["x", "y"]
    model = est.get_model()
    if results["backend"] == "sklearn-hist-gbt":
        pred = model.predict(np.array([[1.0, 1.0], [5.0, 5.0]], dtype=np.float32))
        assert abs(pred[0] - 12.0) < 3.0 and abs(pred[1] - 40.0) < 3.0
        assert results["history"]["valid-score"] > 0.9


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
