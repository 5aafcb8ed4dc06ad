"""``train.torch``: device placement, model and data-loader preparation, ``TorchTrainer``.

Mirrors ``ray.train.torch`` as called by the reference: ``get_device()``
(``/root/reference/my_ray_module.py:124``), ``prepare_data_loader()`` (``:128-129``),
``prepare_model()`` (``:135``), ``TorchTrainer`` (``:244``).
"""
from __future__ import annotations

from typing import Optional

import torch
import torch.distributed as dist
from torch import nn

from ..session import get_context
from ..trainer import TorchTrainer

__all__ = ["TorchTrainer", "get_device", "get_devices", "prepare_model", "prepare_data_loader", "prepare_optimizer",
           "enable_reproducibility"]


def get_device() -> torch.device:
    """``cuda:<local_rank>`` inside a GPU worker, else ``cpu``."""
    ctx = get_context()
    if getattr(ctx, "_use_gpu", False) and torch.cuda.is_available():
        return torch.device("cuda", ctx.get_local_rank())
    return torch.device("cpu")


def get_devices():
    return [get_device()]


def prepare_model(model: nn.Module, *, move_to_device: bool = True, parallel_strategy: Optional[str] = "ddp",
                  fuse: bool = True, ddp_kwargs: Optional[dict] = None) -> nn.Module:
    """Move ``model`` to this worker's device and make it data parallel.

    * ``world_size == 1``: the model is returned unwrapped (state_dict keys carry no prefix), exactly
      like Ray's helper — this is why the reference strips ``module.`` defensively on load.
    * ``world_size > 1``: wrapped in :class:`parallel.DistributedDataParallel` (flat arenas, bucketed
      NVLink all-reduce overlapped with backward, rank-0 broadcast at wrap time). Keys gain ``module.``.
    * ``fuse=True`` (GPU only): ``Linear -> ReLU [-> Dropout]`` chains inside ``nn.Sequential`` containers
      are replaced by modules backed by the fused tcgen05 kernels (same parameters, same keys).
    """
    device = get_device()
    if move_to_device:
        model = model.to(device)
    if fuse and device.type == "cuda":
        from ...models.fused import fuse_linear_chains
        model = fuse_linear_chains(model)
    world = get_context().get_world_size()
    if parallel_strategy is None or world <= 1 or not dist.is_initialized():
        return model
    if parallel_strategy not in ("ddp",):
        raise ValueError(f"parallel_strategy={parallel_strategy!r} is not supported (the reference uses DDP only)")
    from ...parallel.ddp import DistributedDataParallel
    return DistributedDataParallel(model, device=device, **(ddp_kwargs or {}))


def prepare_data_loader(data_loader, *, add_dist_sampler: bool = True, move_to_device: bool = True,
                        auto_transfer: bool = True):
    """Shard a ``DataLoader`` across workers and move its batches to the worker's device.

    With more than one worker the loader is rebuilt around a ``DistributedSampler`` (shuffling iff
    the original sampler was random); ``loader.sampler.set_epoch`` stays available
    (``/root/reference/my_ray_module.py:151``). Batches are copied host->device through pinned memory
    on a side stream, one batch ahead of the consumer.
    """
    from ...data.loader import DeviceLoader, shard_data_loader
    ctx = get_context()
    world, rank = ctx.get_world_size(), ctx.get_world_rank()
    if add_dist_sampler and world > 1:
        data_loader = shard_data_loader(data_loader, world, rank)
    if move_to_device:
        return DeviceLoader(data_loader, get_device(), prefetch=auto_transfer)
    return data_loader


def prepare_optimizer(optimizer):
    """Kept for API parity with Ray (a no-op without AMP)."""
    return optimizer


def enable_reproducibility(seed: int = 0):
    import random
    import numpy as np
    torch.manual_seed(seed)
    random.seed(seed)
    np.random.seed(seed)
    torch.use_deterministic_algorithms(True, warn_only=True)
