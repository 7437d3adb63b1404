"""PyTorch estimator (reference ``python/raydp/torch/__init__.py``)."""
from raydp_b200.torch.estimator import TorchEstimator  # noqa: F401
from raydp_b200.torch.torch_metrics import TorchMetric  # noqa: F401

__all__ = ["TorchEstimator", "TorchMetric"]
