"""``raydp_b200.train.torch`` -- the names scripts reach through ``ray.train.torch`` (the reference's estimator calls
``train.torch.prepare_model``, ``python/raydp/torch/estimator.py:243``; user loops add ``get_device`` /
``prepare_data_loader``)."""
from raydp_b200.train.checkpoint import TorchCheckpoint  # noqa: F401
from raydp_b200.train.config import TorchConfig  # noqa: F401
from raydp_b200.train.torch_utils import get_device, prepare_data_loader, prepare_model, unwrap_model  # noqa: F401
from raydp_b200.train.trainer import TorchTrainer  # noqa: F401
