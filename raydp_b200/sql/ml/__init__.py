"""``pyspark.ml`` stand-in: pipelines, feature stages, linear models, evaluators (see the sub-modules)."""
from raydp_b200.sql.ml.base import Estimator, Model, Pipeline, PipelineModel, Transformer  # noqa: F401
