"""``pyspark.ml`` building blocks: params, ``Transformer`` / ``Estimator`` / ``Pipeline``.

Programs that feed RayDP estimators usually prepare features with ``pyspark.ml.feature`` stages
(``StringIndexer``, ``VectorAssembler``, scalers) chained in a ``Pipeline`` before calling
``fit_on_spark``.  The stages here run on the built-in DataFrame engine; vectors are plain
``array<double>`` columns (Arrow ``list<float64>``), which is also what the dataset / tensor bridge
consumes directly.
"""
from __future__ import annotations

import copy
from typing import Any, Dict, List, Optional


class Params:
    """Keyword parameters with ``getX`` / ``setX`` accessors generated from ``_defaults``."""

    _defaults: Dict[str, Any] = {}

    def __init__(self, **kwargs):
        self._values: Dict[str, Any] = {}
        self._set(**kwargs)

    def _set(self, **kwargs) -> "Params":
        known = self._all_defaults()
        for k, v in kwargs.items():
            if k not in known:
                raise TypeError(f"{type(self).__name__} has no param {k!r} (known: {sorted(known)})")
            if v is not None:
                self._values[k] = v
        return self

    setParams = _set

    @classmethod
    def _all_defaults(cls) -> Dict[str, Any]:
        merged: Dict[str, Any] = {}
        for klass in reversed(cls.__mro__):
            merged.update(getattr(klass, "_defaults", {}))
        return merged

    def getOrDefault(self, name: str):
        if name in self._values:
            return self._values[name]
        return self._all_defaults()[name]

    def isSet(self, name: str) -> bool:
        return name in self._values

    def isDefined(self, name: str) -> bool:
        return name in self._values or self._all_defaults().get(name) is not None

    def hasParam(self, name: str) -> bool:
        return name in self._all_defaults()

    def extractParamMap(self) -> Dict[str, Any]:
        return {**self._all_defaults(), **self._values}

    def explainParams(self) -> str:
        return "\n".join(f"{k}: (default: {d!r}, current: {self.getOrDefault(k)!r})"
                         for k, d in sorted(self._all_defaults().items()))

    def copy(self, extra: Optional[Dict[str, Any]] = None):
        clone = copy.copy(self)
        clone._values = dict(self._values)
        if extra:
            clone._set(**{(k if isinstance(k, str) else k.name): v for k, v in extra.items()})
        return clone

    def __getattr__(self, item: str):
        # getInputCol() / setInputCol(v) for every declared param
        if item.startswith(("get", "set")) and len(item) > 3:
            name = item[3].lower() + item[4:]
            if name in type(self)._all_defaults():
                if item.startswith("get"):
                    return lambda: self.getOrDefault(name)
                return lambda value: self._set(**{name: value})
        raise AttributeError(f"{type(self).__name__!r} object has no attribute {item!r}")

    def __repr__(self):
        shown = ", ".join(f"{k}={v!r}" for k, v in self._values.items())
        return f"{type(self).__name__}({shown})"


class Transformer(Params):
    def transform(self, dataset, params: Optional[Dict[str, Any]] = None):
        stage = self.copy(params) if params else self
        return stage._transform(dataset)

    def _transform(self, dataset):
        raise NotImplementedError


class Estimator(Params):
    def fit(self, dataset, params: Optional[Dict[str, Any]] = None):
        stage = self.copy(params) if params else self
        return stage._fit(dataset)

    def _fit(self, dataset):
        raise NotImplementedError


class Model(Transformer):
    pass


class Pipeline(Estimator):
    _defaults = {"stages": None}

    def __init__(self, stages: Optional[List[Any]] = None):
        super().__init__(stages=stages)

    def getStages(self) -> List[Any]:
        return list(self.getOrDefault("stages") or [])

    def setStages(self, value: List[Any]) -> "Pipeline":
        return self._set(stages=value)

    def _fit(self, dataset) -> "PipelineModel":
        fitted = []
        stages = self.getStages()
        last_estimator = max((i for i, s in enumerate(stages) if isinstance(s, Estimator)), default=-1)
        current = dataset
        for i, stage in enumerate(stages):
            if isinstance(stage, Estimator):
                model = stage.fit(current)
                fitted.append(model)
                if i < last_estimator:
                    current = model.transform(current)
            elif isinstance(stage, Transformer):
                fitted.append(stage)
                if i < last_estimator:
                    current = stage.transform(current)
            else:
                raise TypeError(f"Cannot recognize a pipeline stage of type {type(stage)}.")
        return PipelineModel(fitted)


class PipelineModel(Model):
    def __init__(self, stages: List[Transformer]):
        super().__init__()
        self.stages = list(stages)

    def _transform(self, dataset):
        for stage in self.stages:
            dataset = stage.transform(dataset)
        return dataset

    def save(self, path: str) -> None:
        import os
        import cloudpickle
        os.makedirs(path, exist_ok=True)
        with open(os.path.join(path, "pipeline.pkl"), "wb") as fh:
            cloudpickle.dump(self, fh)

    def write(self):
        return _Writer(self)

    @classmethod
    def load(cls, path: str) -> "PipelineModel":
        import os
        import pickle
        with open(os.path.join(path, "pipeline.pkl"), "rb") as fh:
            return pickle.load(fh)


class _Writer:
    def __init__(self, model):
        self._model = model

    def overwrite(self) -> "_Writer":
        return self

    def save(self, path: str) -> None:
        self._model.save(path)
