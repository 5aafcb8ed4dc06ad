"""Trainer API with the surface of ``ray.train`` as used by the reference
(``/root/reference/my_ray_module.py:16-25``): ``ScalingConfig``, ``RunConfig``, ``Checkpoint``,
``CheckpointConfig``, ``report``, ``get_context``, ``get_checkpoint``; ``train.torch`` holds
``TorchTrainer``, ``get_device``, ``prepare_model``, ``prepare_data_loader``."""
from .checkpoint import Checkpoint
from .config import CheckpointConfig, FailureConfig, RunConfig, ScalingConfig
from .result import Result
from .session import TrainContext, get_checkpoint, get_context, report
from .trainer import TorchTrainer, TrainingFailedError

__all__ = ["Checkpoint", "CheckpointConfig", "FailureConfig", "RunConfig", "ScalingConfig", "Result",
           "TrainContext", "get_checkpoint", "get_context", "report", "TorchTrainer", "TrainingFailedError"]
