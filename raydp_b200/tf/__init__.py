from raydp_b200.tf.estimator import TFEstimator  # noqa: F401

__all__ = ["TFEstimator"]
