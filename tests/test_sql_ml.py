"""``pyspark.ml`` stand-in: feature stages, pipelines, linear models, evaluators, model selection."""
import numpy as np
import pytest

from raydp_b200.sql.ml import Pipeline, PipelineModel
from raydp_b200.sql.ml.feature import (Bucketizer, HashingTF, Imputer, MinMaxScaler, Normalizer, OneHotEncoder,
                                       QuantileDiscretizer, SQLTransformer, StandardScaler, StopWordsRemover,
                                       StringIndexer, Tokenizer, VectorAssembler)
from raydp_b200.sql.ml.linalg import Vectors
from raydp_b200.sql.ml.models import (BinaryClassificationEvaluator, CrossValidator, KMeans, LinearRegression,
                                      LogisticRegression, MulticlassClassificationEvaluator, ParamGridBuilder,
                                      RegressionEvaluator, TrainValidationSplit)


@pytest.fixture(scope="module")
def frame(spark_session):
    return spark_session.createDataFrame(
        [("a", 1.0, 10.0, "Hi there you"), ("b", 2.0, None, "the cat"), ("a", 3.0, 30.0, "A b"), ("c", 4.0, 40.0, None)],
        ["k", "v", "x", "t"])


def test_feature_pipeline(frame):
    indexer = StringIndexer(inputCol="k", outputCol="ki").fit(frame)
    assert indexer.labels == ["a", "b", "c"] and [r.ki for r in indexer.transform(frame).collect()] == [0.0, 1.0, 0.0, 2.0]
    pipe = Pipeline(stages=[Imputer(inputCols=["x"], outputCols=["xf"]), StringIndexer(inputCol="k", outputCol="ki"),
                            OneHotEncoder(inputCol="ki", outputCol="oh"),
                            VectorAssembler(inputCols=["v", "xf", "oh"], outputCol="f"),
                            StandardScaler(inputCol="f", outputCol="fs", withMean=True),
                            MinMaxScaler(inputCol="f", outputCol="fm")])
    rows = pipe.fit(frame).transform(frame).collect()
    assert abs(rows[1].xf - 80.0 / 3) < 1e-12 and rows[0].oh == [1.0, 0.0] and rows[3].oh == [0.0, 0.0]
    assert rows[0].f == [1.0, 10.0, 1.0, 0.0] and rows[0].fm == [0.0, 0.0, 1.0, 0.0] and rows[3].fm[:2] == [1.0, 1.0]
    scaled = np.asarray([r.fs for r in rows])
    assert np.allclose(scaled.mean(axis=0), 0.0) and np.allclose(scaled.std(axis=0, ddof=1), 1.0)
    with pytest.raises(Exception):
        StringIndexer(inputCol="k", outputCol="ki").fit(frame.filter("k != 'c'")).transform(frame).collect()
    kept = StringIndexer(inputCol="k", outputCol="ki", handleInvalid="keep").fit(frame.filter("k != 'c'")).transform(frame)
    assert [r.ki for r in kept.collect()] == [0.0, 1.0, 0.0, 2.0]
    assert Vectors.dense(1, 2).dot([3, 4]) == 11.0 and Vectors.sparse(3, {1: 2.0}).toArray().tolist() == [0.0, 2.0, 0.0]


def test_numeric_and_text_stages(frame):
    assert [r.b for r in Bucketizer(splits=[0, 2, 3, 10], inputCol="v", outputCol="b").transform(frame).collect()] == \
        [0.0, 1.0, 2.0, 2.0]
    disc = QuantileDiscretizer(numBuckets=2, inputCol="v", outputCol="q").fit(frame).transform(frame)
    assert [r.q for r in disc.collect()] == [0.0, 1.0, 1.0, 1.0]
    words = Tokenizer(inputCol="t", outputCol="w").transform(frame.filter("t is not null"))
    assert [r.w2 for r in StopWordsRemover(inputCol="w", outputCol="w2").transform(words).collect()] == \
        [["hi"], ["cat"], ["b"]]
    tf = HashingTF(inputCol="w", outputCol="tf", numFeatures=8).transform(words).collect()
    assert [sum(r.tf) for r in tf] == [3.0, 2.0, 2.0]
    out = SQLTransformer(statement="SELECT k, v * 2 AS v2 FROM __THIS__ WHERE v > 1").transform(frame).collect()
    assert [(r.k, r.v2) for r in out] == [("b", 4.0), ("a", 6.0), ("c", 8.0)]
    vec = VectorAssembler(inputCols=["v", "v"], outputCol="f").transform(frame)
    unit = Normalizer(inputCol="f", outputCol="n").transform(vec).collect()
    assert np.allclose([np.linalg.norm(r.n) for r in unit], 1.0)


def test_models_evaluators_and_selection(spark_session, tmp_path):
    rng = np.random.default_rng(0)
    x = rng.normal(size=(400, 3))
    y = x @ np.array([1.5, -2.0, 0.5]) + 3.0 + rng.normal(scale=0.01, size=400)
    cls = (x[:, 0] - x[:, 1] > 0).astype(float)
    df = spark_session.createDataFrame([(float(a), float(b), float(c), float(t), float(k))
                                        for (a, b, c), t, k in zip(x, y, cls)], ["a", "b", "c", "y", "cls"])
    feats = VectorAssembler(inputCols=["a", "b", "c"], outputCol="features").transform(df)
    lr = LinearRegression(labelCol="y").fit(feats)
    assert np.allclose(lr.coefficients.toArray(), [1.5, -2.0, 0.5], atol=5e-3) and abs(lr.intercept - 3.0) < 5e-3
    assert lr.evaluate(feats).r2 > 0.9999 and RegressionEvaluator(labelCol="y").evaluate(lr.transform(feats)) < 0.02
    pred = LogisticRegression(labelCol="cls", regParam=0.01).fit(feats).transform(feats)
    assert MulticlassClassificationEvaluator(labelCol="cls", metricName="accuracy").evaluate(pred) > 0.97
    assert BinaryClassificationEvaluator(labelCol="cls", rawPredictionCol="probability").evaluate(pred) > 0.99
    clusters = KMeans(k=2, seed=1).fit(feats).transform(feats).groupBy("prediction").count().collect()
    assert sorted(r.prediction for r in clusters) == [0, 1] and sum(r["count"] for r in clusters) == 400
    grid = ParamGridBuilder().addGrid("regParam", [0.0, 10.0]).build()
    cv = CrossValidator(estimator=LinearRegression(labelCol="y"), estimatorParamMaps=grid,
                        evaluator=RegressionEvaluator(labelCol="y"), numFolds=3).fit(feats)
    assert cv.avgMetrics[0] < cv.avgMetrics[1] and cv.bestModel.getOrDefault("regParam") == 0.0
    tvs = TrainValidationSplit(estimator=LinearRegression(labelCol="y"), estimatorParamMaps=grid,
                               evaluator=RegressionEvaluator(labelCol="y", metricName="r2")).fit(feats)
    assert tvs.validationMetrics[0] > tvs.validationMetrics[1]
    pipe = Pipeline(stages=[VectorAssembler(inputCols=["a", "b", "c"], outputCol="raw"),
                            StandardScaler(inputCol="raw", outputCol="features", withMean=True),
                            LinearRegression(labelCol="y")]).fit(df)
    pipe.write().overwrite().save(str(tmp_path / "m"))
    again = PipelineModel.load(str(tmp_path / "m")).transform(df)
    assert RegressionEvaluator(labelCol="y", metricName="r2").evaluate(again) > 0.9999


def test_pyspark_ml_import_names():
    import subprocess
    import sys
    code = ("import raydp; from pyspark.ml import Pipeline; from pyspark.ml.feature import VectorAssembler, StringIndexer; "
            "from pyspark.ml.regression import LinearRegression; from pyspark.ml.evaluation import RegressionEvaluator; "
            "from pyspark.ml.linalg import Vectors; from pyspark.ml.tuning import CrossValidator, ParamGridBuilder; print('ok')")
    import os
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    proc = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True,
                          env=dict(os.environ, PYTHONPATH=root + os.pathsep + os.environ.get("PYTHONPATH", "")))
    assert proc.returncode == 0 and "ok" in proc.stdout, proc.stderr[-1500:]


def test_assembled_vector_column_feeds_the_torch_bridge(ray_cluster):
    """``VectorAssembler`` output (``array<double>``) is a valid feature column of ``Dataset.to_torch``: it
    contributes its width to the feature matrix."""
    import torch
    from raydp_b200 import data
    ds = data.from_items([{"f": [float(i), float(2 * i)], "s": float(i) + 0.5, "y": float(i)} for i in range(6)])
    batches = list(ds.to_torch(label_column="y", feature_columns=["s", "f"], feature_column_dtypes=torch.float,
                               label_column_dtype=torch.float, batch_size=4))
    feats, label = batches[0]
    assert feats.shape == (4, 3) and feats[1].tolist() == [1.5, 1.0, 2.0] and label.shape == (4, 1)
    per_col = list(ds.to_torch(label_column="y", feature_columns=["s", "f"],
                               feature_column_dtypes=[torch.float, torch.float], label_column_dtype=torch.float,
                               batch_size=6))[0][0]
    assert per_col[0].shape == (6, 1) and per_col[1].shape == (6, 2)
