"""Training runtime: worker group, sessions, checkpoints (the Ray Train surface RayDP uses)."""
from raydp_b200.train.checkpoint import Checkpoint, TorchCheckpoint  # noqa: F401
from raydp_b200.train.config import (CheckpointConfig, FailureConfig, RunConfig,  # noqa: F401
                                     ScalingConfig, TorchConfig)
from raydp_b200.train.session import (get_checkpoint, get_context, get_dataset_shard,  # noqa: F401
                                      report)
from raydp_b200.train.torch_utils import get_device, prepare_data_loader, prepare_model  # noqa: F401
from raydp_b200.train.trainer import Result, TorchTrainer, TrainingFailedError  # noqa: F401
from raydp_b200.train.legacy import Trainer, TrainingCallback  # noqa: F401,E402
from raydp_b200.train import torch  # noqa: F401,E402  (train.torch.prepare_model, as in ray.train.torch)
