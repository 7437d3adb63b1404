"""Linear models, k-means, evaluators and model selection (``pyspark.ml.regression`` / ``classification`` /
``clustering`` / ``evaluation`` / ``tuning``).

Fitting gathers per-partition sufficient statistics where the model allows it (the normal equations of
linear regression need only ``X^T X`` and ``X^T y``); the iterative models (logistic regression,
k-means) work on the assembled feature matrix with torch / numpy.  Predictions are Arrow kernels in
the lazy plan, like the feature stages.
"""
from __future__ import annotations

import itertools
from typing import Any, Dict, List, Optional

import numpy as np
import pyarrow as pa
import pyarrow.compute as pc

from raydp_b200.sql.ml.base import Estimator, Model, Params
from raydp_b200.sql.ml.feature import _arrow_expr, _col, _from_matrix, _to_matrix
from raydp_b200.sql.ml.linalg import DenseVector


def _collect_xy(dataset, features: str, label: Optional[str] = None, weight: Optional[str] = None):
    cols = [features] + ([label] if label else []) + ([weight] if weight else [])
    table = dataset.select(*cols)._collect_table()
    x = _to_matrix(table.column(features).combine_chunks())
    y = pc.cast(table.column(label), pa.float64()).to_numpy() if label else None
    w = pc.cast(table.column(weight), pa.float64()).to_numpy() if weight else None
    return x, y, w


class _Predictor(Estimator):
    _defaults = {"featuresCol": "features", "labelCol": "label", "predictionCol": "prediction", "weightCol": None}


# ------------------------------------------------------------------------------------------ regression
class LinearRegression(_Predictor):
    _defaults = {"regParam": 0.0, "elasticNetParam": 0.0, "fitIntercept": True, "maxIter": 100, "tol": 1e-6,
                 "standardization": True, "solver": "auto"}

    def __init__(self, **kwargs):
        super().__init__(**kwargs)

    def _fit(self, dataset) -> "LinearRegressionModel":
        features, label = self.getOrDefault("featuresCol"), self.getOrDefault("labelCol")
        intercept = bool(self.getOrDefault("fitIntercept"))

        def gram(rows):
            import numpy as _np
            data = [(r[0], r[1]) for r in rows if r[0] is not None and r[1] is not None]
            if not data:
                return []
            x = _np.asarray([d[0] for d in data], dtype=_np.float64)
            y = _np.asarray([d[1] for d in data], dtype=_np.float64)
            if intercept:
                x = _np.hstack([x, _np.ones((len(x), 1))])
            return [(x.T @ x, x.T @ y, float(y @ y), len(y))]
        parts = dataset.select(features, label).rdd.mapPartitions(gram).collect()
        if not parts:
            raise ValueError("LinearRegression: no rows to fit on")
        xtx = np.sum([p[0] for p in parts], axis=0)
        xty = np.sum([p[1] for p in parts], axis=0)
        n = sum(p[3] for p in parts)
        reg = float(self.getOrDefault("regParam")) * n
        penalty = np.eye(len(xtx)) * reg
        if intercept:
            penalty[-1, -1] = 0.0
        beta = np.linalg.lstsq(xtx + penalty, xty, rcond=None)[0]
        coef, b = (beta[:-1], float(beta[-1])) if intercept else (beta, 0.0)
        return LinearRegressionModel(coef, b)._set(**self._values)


class LinearRegressionModel(Model):
    _defaults = dict(_Predictor._defaults, **LinearRegression._defaults)

    def __init__(self, coefficients: np.ndarray, intercept: float):
        super().__init__()
        self.coefficients, self.intercept = DenseVector(coefficients), float(intercept)

    @property
    def numFeatures(self) -> int:
        return len(self.coefficients)

    def predict(self, features) -> float:
        return float(np.dot(self.coefficients.toArray(), np.asarray(list(features), dtype=np.float64)) + self.intercept)

    def _transform(self, dataset):
        w, b = self.coefficients.toArray(), self.intercept
        return dataset.withColumn(self.getOrDefault("predictionCol"), _arrow_expr(
            lambda a: pa.array(_to_matrix(a) @ w + b), self.getOrDefault("featuresCol"), name="linear_predict",
            empty=pa.float64()))

    def evaluate(self, dataset) -> "RegressionSummary":
        return RegressionSummary(self.transform(dataset), self.getOrDefault("labelCol"), self.getOrDefault("predictionCol"))

    @property
    def summary(self):
        raise RuntimeError("call model.evaluate(dataset) for a summary on a dataset")


class RegressionSummary:
    def __init__(self, predictions, label: str, prediction: str):
        ev = RegressionEvaluator(labelCol=label, predictionCol=prediction)
        self.predictions = predictions
        self.rootMeanSquaredError = ev.evaluate(predictions, {"metricName": "rmse"})
        self.meanSquaredError = ev.evaluate(predictions, {"metricName": "mse"})
        self.meanAbsoluteError = ev.evaluate(predictions, {"metricName": "mae"})
        self.r2 = ev.evaluate(predictions, {"metricName": "r2"})


# ------------------------------------------------------------------------------------------ classification
class LogisticRegression(_Predictor):
    _defaults = {"regParam": 0.0, "elasticNetParam": 0.0, "fitIntercept": True, "maxIter": 100, "tol": 1e-6,
                 "threshold": 0.5, "family": "auto", "probabilityCol": "probability", "rawPredictionCol": "rawPrediction",
                 "standardization": True}

    def __init__(self, **kwargs):
        super().__init__(**kwargs)

    def _fit(self, dataset) -> "LogisticRegressionModel":
        import torch
        x, y, w = _collect_xy(dataset, self.getOrDefault("featuresCol"), self.getOrDefault("labelCol"),
                              self.getOrDefault("weightCol"))
        classes = int(y.max()) + 1 if len(y) else 2
        classes = max(classes, 2)
        mean, std = x.mean(axis=0), x.std(axis=0)
        std = np.where(std > 0, std, 1.0)
        xs = torch.tensor((x - mean) / std, dtype=torch.float64)
        yt = torch.tensor(y, dtype=torch.long)
        wt = torch.tensor(w if w is not None else np.ones(len(y)), dtype=torch.float64)
        outputs = 1 if classes == 2 and self.getOrDefault("family") != "multinomial" else classes
        weight = torch.zeros((xs.shape[1], outputs), dtype=torch.float64, requires_grad=True)
        bias = torch.zeros(outputs, dtype=torch.float64, requires_grad=bool(self.getOrDefault("fitIntercept")))
        reg = float(self.getOrDefault("regParam"))
        opt = torch.optim.LBFGS([weight, bias] if bias.requires_grad else [weight], max_iter=int(self.getOrDefault("maxIter")),
                                tolerance_grad=float(self.getOrDefault("tol")), line_search_fn="strong_wolfe")

        def closure():
            opt.zero_grad()
            logits = xs @ weight + bias
            if outputs == 1:
                loss = torch.nn.functional.binary_cross_entropy_with_logits(logits[:, 0], yt.double(), weight=wt, reduction="sum")
            else:
                loss = (torch.nn.functional.cross_entropy(logits, yt, reduction="none") * wt).sum()
            loss = loss / wt.sum() + 0.5 * reg * (weight ** 2).sum()
            loss.backward()
            return loss
        opt.step(closure)
        wn = weight.detach().numpy() / std[:, None]                      # back to the original feature scale
        bn = bias.detach().numpy() - (mean[:, None] * wn).sum(axis=0)
        return LogisticRegressionModel(wn, bn, classes)._set(**self._values)


class LogisticRegressionModel(Model):
    _defaults = dict(_Predictor._defaults, **LogisticRegression._defaults)

    def __init__(self, weights: np.ndarray, bias: np.ndarray, num_classes: int):
        super().__init__()
        self.coefficientMatrix, self.interceptVector, self.numClasses = weights, bias, num_classes

    @property
    def coefficients(self) -> DenseVector:
        return DenseVector(self.coefficientMatrix[:, 0])

    @property
    def intercept(self) -> float:
        return float(self.interceptVector[0])

    def _probabilities(self, m: np.ndarray) -> np.ndarray:
        logits = m @ self.coefficientMatrix + self.interceptVector
        if logits.shape[1] == 1:
            p1 = 1.0 / (1.0 + np.exp(-logits[:, 0]))
            return np.stack([1.0 - p1, p1], axis=1)
        logits = logits - logits.max(axis=1, keepdims=True)
        e = np.exp(logits)
        return e / e.sum(axis=1, keepdims=True)

    def _transform(self, dataset):
        threshold = float(self.getOrDefault("threshold"))
        features = self.getOrDefault("featuresCol")

        def proba(a):
            return _from_matrix(self._probabilities(_to_matrix(a)))

        def predict(a):
            p = self._probabilities(_to_matrix(a))
            if p.shape[1] == 2:
                return pa.array((p[:, 1] > threshold).astype(np.float64))
            return pa.array(p.argmax(axis=1).astype(np.float64))
        out = dataset.withColumn(self.getOrDefault("probabilityCol"), _arrow_expr(proba, features, name="lr_probability"))
        return out.withColumn(self.getOrDefault("predictionCol"),
                              _arrow_expr(predict, features, name="lr_predict", empty=pa.float64()))


# ------------------------------------------------------------------------------------------ clustering
class KMeans(Estimator):
    _defaults = {"featuresCol": "features", "predictionCol": "prediction", "k": 2, "maxIter": 20, "seed": 0, "tol": 1e-4,
                 "initMode": "k-means||"}

    def __init__(self, **kwargs):
        super().__init__(**kwargs)

    def _fit(self, dataset) -> "KMeansModel":
        x, _, _ = _collect_xy(dataset, self.getOrDefault("featuresCol"))
        k = int(self.getOrDefault("k"))
        rng = np.random.default_rng(int(self.getOrDefault("seed")))
        centers = [x[rng.integers(len(x))]]
        for _ in range(1, k):                                            # k-means++ seeding
            d2 = np.min([((x - c) ** 2).sum(axis=1) for c in centers], axis=0)
            centers.append(x[rng.choice(len(x), p=d2 / d2.sum())] if d2.sum() > 0 else x[rng.integers(len(x))])
        centers = np.asarray(centers)
        for _ in range(int(self.getOrDefault("maxIter"))):
            assign = ((x[:, None, :] - centers[None, :, :]) ** 2).sum(axis=2).argmin(axis=1)
            new = np.asarray([x[assign == j].mean(axis=0) if (assign == j).any() else centers[j] for j in range(k)])
            shift = np.sqrt(((new - centers) ** 2).sum(axis=1)).max()
            centers = new
            if shift < float(self.getOrDefault("tol")):
                break
        return KMeansModel(centers)._set(**self._values)


class KMeansModel(Model):
    _defaults = dict(KMeans._defaults)

    def __init__(self, centers: np.ndarray):
        super().__init__()
        self._centers = np.asarray(centers)

    def clusterCenters(self) -> List[np.ndarray]:
        return [c for c in self._centers]

    def _transform(self, dataset):
        centers = self._centers
        return dataset.withColumn(self.getOrDefault("predictionCol"), _arrow_expr(
            lambda a: pa.array(((_to_matrix(a)[:, None, :] - centers[None, :, :]) ** 2).sum(axis=2).argmin(axis=1).astype(np.int32)),
            self.getOrDefault("featuresCol"), name="kmeans_predict", empty=pa.int32()))


# ------------------------------------------------------------------------------------------ evaluation
class Evaluator(Params):
    def evaluate(self, dataset, params: Optional[Dict[str, Any]] = None) -> float:
        ev = self.copy(params) if params else self
        return ev._evaluate(dataset)

    def isLargerBetter(self) -> bool:
        return True


class RegressionEvaluator(Evaluator):
    _defaults = {"predictionCol": "prediction", "labelCol": "label", "metricName": "rmse"}

    def __init__(self, **kwargs):
        super().__init__(**kwargs)

    def _evaluate(self, dataset) -> float:
        t = dataset.select(self.getOrDefault("labelCol"), self.getOrDefault("predictionCol"))._collect_table()
        y, p = (pc.cast(t.column(i), pa.float64()).to_numpy() for i in (0, 1))
        err = y - p
        metric = self.getOrDefault("metricName")
        if metric == "mse":
            return float(np.mean(err ** 2))
        if metric == "rmse":
            return float(np.sqrt(np.mean(err ** 2)))
        if metric == "mae":
            return float(np.mean(np.abs(err)))
        if metric == "r2":
            return float(1.0 - (err ** 2).sum() / ((y - y.mean()) ** 2).sum())
        raise ValueError(f"unknown regression metric {metric!r}")

    def isLargerBetter(self) -> bool:
        return self.getOrDefault("metricName") == "r2"


class BinaryClassificationEvaluator(Evaluator):
    _defaults = {"rawPredictionCol": "rawPrediction", "labelCol": "label", "metricName": "areaUnderROC"}

    def __init__(self, **kwargs):
        super().__init__(**kwargs)

    def _evaluate(self, dataset) -> float:
        raw = self.getOrDefault("rawPredictionCol")
        if raw not in dataset.columns and "probability" in dataset.columns:
            raw = "probability"
        t = dataset.select(self.getOrDefault("labelCol"), raw)._collect_table()
        y = pc.cast(t.column(0), pa.float64()).to_numpy()
        scores = _to_matrix(t.column(1).combine_chunks())
        s = scores[:, -1]
        order = np.argsort(-s, kind="stable")
        y_sorted = y[order]
        tp = np.concatenate([[0], np.cumsum(y_sorted)])
        fp = np.concatenate([[0], np.cumsum(1 - y_sorted)])
        distinct = np.concatenate([np.flatnonzero(np.diff(s[order])) + 1, [len(s)]])
        tp, fp = np.concatenate([[0], tp[distinct]]), np.concatenate([[0], fp[distinct]])
        pos, neg = max(y.sum(), 1e-12), max((1 - y).sum(), 1e-12)
        if self.getOrDefault("metricName") == "areaUnderROC":
            return float(np.trapezoid(tp / pos, fp / neg))
        precision = np.where(tp + fp > 0, tp / np.maximum(tp + fp, 1e-12), 1.0)
        return float(np.trapezoid(precision, tp / pos))


class MulticlassClassificationEvaluator(Evaluator):
    _defaults = {"predictionCol": "prediction", "labelCol": "label", "metricName": "f1"}

    def __init__(self, **kwargs):
        super().__init__(**kwargs)

    def _evaluate(self, dataset) -> float:
        t = dataset.select(self.getOrDefault("labelCol"), self.getOrDefault("predictionCol"))._collect_table()
        y, p = (pc.cast(t.column(i), pa.float64()).to_numpy() for i in (0, 1))
        metric = self.getOrDefault("metricName")
        if metric == "accuracy":
            return float((y == p).mean())
        labels, counts = np.unique(y, return_counts=True)
        prec, rec = [], []
        for lab in labels:
            tp = float(((p == lab) & (y == lab)).sum())
            prec.append(tp / max((p == lab).sum(), 1e-12) if (p == lab).any() else 0.0)
            rec.append(tp / max((y == lab).sum(), 1e-12))
        prec, rec, wts = np.asarray(prec), np.asarray(rec), counts / counts.sum()
        if metric == "weightedPrecision":
            return float((prec * wts).sum())
        if metric == "weightedRecall":
            return float((rec * wts).sum())
        if metric == "f1":
            f1 = np.where(prec + rec > 0, 2 * prec * rec / np.maximum(prec + rec, 1e-12), 0.0)
            return float((f1 * wts).sum())
        raise ValueError(f"unknown multiclass metric {metric!r}")


# ------------------------------------------------------------------------------------------ model selection
class ParamGridBuilder:
    def __init__(self):
        self._grid: Dict[Any, List[Any]] = {}

    def addGrid(self, param, values) -> "ParamGridBuilder":
        self._grid[param if isinstance(param, str) else getattr(param, "name", param)] = list(values)
        return self

    def baseOn(self, *args) -> "ParamGridBuilder":
        pairs = args[0].items() if len(args) == 1 and isinstance(args[0], dict) else args
        for k, v in pairs:
            self.addGrid(k, [v])
        return self

    def build(self) -> List[Dict[str, Any]]:
        keys = list(self._grid)
        return [dict(zip(keys, combo)) for combo in itertools.product(*[self._grid[k] for k in keys])]


class _Validator(Estimator):
    _defaults = {"estimator": None, "estimatorParamMaps": None, "evaluator": None, "seed": 0, "parallelism": 1}

    def _score(self, train, test, params) -> float:
        est, ev = self.getOrDefault("estimator"), self.getOrDefault("evaluator")
        return ev.evaluate(est.fit(train, params).transform(test))

    def _pick(self, scores: List[float]) -> int:
        ev = self.getOrDefault("evaluator")
        return int(np.argmax(scores) if ev.isLargerBetter() else np.argmin(scores))


class CrossValidator(_Validator):
    _defaults = {"numFolds": 3}

    def __init__(self, **kwargs):
        super().__init__(**kwargs)

    def _fit(self, dataset) -> "CrossValidatorModel":
        from raydp_b200.sql import functions as F
        k, seed = int(self.getOrDefault("numFolds")), int(self.getOrDefault("seed"))
        grid = self.getOrDefault("estimatorParamMaps") or [{}]
        folded = dataset.withColumn("__fold", (F.rand(seed) * k).cast("int")).cache()
        scores = []
        for params in grid:
            per_fold = [self._score(folded.filter(F.col("__fold") != i).drop("__fold"),
                                    folded.filter(F.col("__fold") == i).drop("__fold"), params) for i in range(k)]
            scores.append(float(np.mean(per_fold)))
        best = self._pick(scores)
        model = self.getOrDefault("estimator").fit(dataset, grid[best])
        return CrossValidatorModel(model, scores)


class CrossValidatorModel(Model):
    def __init__(self, best_model, avg_metrics: List[float]):
        super().__init__()
        self.bestModel, self.avgMetrics = best_model, avg_metrics

    def _transform(self, dataset):
        return self.bestModel.transform(dataset)


class TrainValidationSplit(_Validator):
    _defaults = {"trainRatio": 0.75}

    def __init__(self, **kwargs):
        super().__init__(**kwargs)

    def _fit(self, dataset) -> "TrainValidationSplitModel":
        ratio = float(self.getOrDefault("trainRatio"))
        train, test = dataset.randomSplit([ratio, 1 - ratio], seed=int(self.getOrDefault("seed")))
        grid = self.getOrDefault("estimatorParamMaps") or [{}]
        scores = [self._score(train, test, params) for params in grid]
        model = self.getOrDefault("estimator").fit(dataset, grid[self._pick(scores)])
        return TrainValidationSplitModel(model, scores)


class TrainValidationSplitModel(Model):
    def __init__(self, best_model, validation_metrics: List[float]):
        super().__init__()
        self.bestModel, self.validationMetrics = best_model, validation_metrics

    def _transform(self, dataset):
        return self.bestModel.transform(dataset)
