from raydp_b200.xgboost.estimator import XGBoostEstimator  # noqa: F401

__all__ = ["XGBoostEstimator"]
