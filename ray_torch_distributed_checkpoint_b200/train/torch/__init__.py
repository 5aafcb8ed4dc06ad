"""``train.torch``: device placement, model and data-loader preparation, ``TorchTrainer``.

Mirrors ``ray.train.torch`` as called by the reference: ``get_device()``
(``/root/reference/my_ray_module.py:124``), ``prepare_data_loader()`` (``:128-129``),
``prepare_model()`` (``:135``), ``TorchTrainer`` (``:244``).
"""
from __future__ import annotations

import os
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
                  fuse=True, ddp_kwargs: Optional[dict] = None, engine_kwargs: Optional[dict] = None) -> nn.Module:
    """Move ``model`` to this worker's device and make it data parallel (ref: ``ray.train.torch.prepare_model`` at
    ``my_ray_module.py:135``).

    On a GPU worker, ``fuse`` picks how the model runs:

    * ``fuse=True`` / ``"engine"`` — when ``model`` has the reference's shape (``Flatten`` + ``Sequential(Linear, ReLU,
      Dropout, Linear, ReLU, Dropout, Linear, ReLU)``, see ``parallel.engine.match_reference_mlp``) the result is a
      ``parallel.EngineModel``: together with ``nn.CrossEntropyLoss`` and ``optim.SGD`` of this package, the reference's
      literal loop (``pred = model(X); loss = loss_fn(pred, y); optimizer.zero_grad(); loss.backward(); optimizer.step()``)
      replays ONE captured CUDA graph of the fused kernels per step, gradient reduction over NVLink and optimizer included.
      Its ``state_dict()`` keys are the bare module's with one worker and ``module.``-prefixed with several, exactly like
      the wrapped-or-not result of Ray's helper. Any other model falls through to the next mode.
    * ``fuse="modules"`` (and the fallback of ``True``) — ``Linear -> ReLU [-> Dropout]`` chains inside ``nn.Sequential``
      containers are replaced by modules backed by the fused tcgen05 kernels under autograd (same parameters, same
      keys); with more than one worker the model is wrapped in :class:`parallel.DistributedDataParallel` (flat arenas,
      bucketed NVLink all-reduce overlapped with backward, rank-0 broadcast at wrap time; keys gain ``module.``).
    * ``fuse=False`` — plain torch modules, same wrapper.

    CPU workers (gloo) always take the last form without fusion.
    """
    device = get_device()
    world = get_context().get_world_size()
    distributed = world > 1 and dist.is_initialized() and parallel_strategy is not None
    if parallel_strategy not in (None, "ddp"):
        raise ValueError(f"parallel_strategy={parallel_strategy!r} is not supported (the reference uses DDP only)")
    if device.type == "cuda" and fuse in (True, "engine") and move_to_device:
        from ...parallel.engine import match_reference_mlp
        if match_reference_mlp(model) is not None:
            from ...parallel.engine_api import EngineModel
            return EngineModel(model, device=device, distributed=distributed, **(engine_kwargs or {}))
        if fuse == "engine":
            raise ValueError("fuse='engine' needs the reference's network shape (parallel.engine.match_reference_mlp)")
    if move_to_device:
        model = model.to(device)
    if fuse and device.type == "cuda":
        from ...models.fused import fuse_linear_chains
        model = fuse_linear_chains(model)
    if not distributed:
        return model
    from ...parallel.ddp import DistributedDataParallel
    return DistributedDataParallel(model, device=device, **(ddp_kwargs or {}))


def _u8_image_source(data_loader):
    """(uint8 images [N, ...], labels [N], mean, std) when ``data_loader`` iterates a map-style image dataset that holds
    its pixels as one uint8 tensor and normalises with a single mean / std (``data.FashionMNISTSynthetic``, torchvision's
    MNIST family with ``ToTensor() + Normalize((m,), (s,))``); else None."""
    import torch.utils.data as tud
    ds = getattr(data_loader, "dataset", None)
    if not isinstance(data_loader, tud.DataLoader) or ds is None or data_loader.batch_size is None:
        return None
    if data_loader.num_workers != 0 or data_loader.collate_fn is not tud.default_collate:
        return None
    if not isinstance(data_loader.sampler, (tud.RandomSampler, tud.SequentialSampler)):
        return None
    imgs, labels = getattr(ds, "data", None), getattr(ds, "targets", None)
    if not isinstance(imgs, torch.Tensor) or imgs.dtype != torch.uint8 or imgs.dim() != 3 or labels is None:
        return None
    labels = torch.as_tensor(labels)
    if labels.dtype != torch.int64 or labels.shape[0] != imgs.shape[0]:
        return None
    from ...data.datasets import FashionMNISTSynthetic
    if isinstance(ds, FashionMNISTSynthetic):
        return None if ds.raw else (imgs, labels, 0.5, 0.5)
    tf = getattr(ds, "transform", None)
    steps = getattr(tf, "transforms", None)
    if getattr(ds, "target_transform", None) is not None or not steps or len(steps) != 2:
        return None
    if type(steps[0]).__name__ != "ToTensor" or type(steps[1]).__name__ != "Normalize":
        return None
    mean, std = steps[1].mean, steps[1].std
    try:
        if len(mean) != 1 or len(std) != 1:
            return None
        return imgs, labels, float(mean[0]), float(std[0])
    except TypeError:
        return imgs, labels, float(mean), float(std)


def prepare_data_loader(data_loader, *, add_dist_sampler: bool = True, move_to_device: bool = True,
                        auto_transfer: bool = True, u8_fast_path: bool = True, transfer: Optional[str] = None):
    """Shard a ``DataLoader`` across workers and move its batches to the worker's device (ref:
    ``ray.train.torch.prepare_data_loader`` at ``my_ray_module.py:128-129``).

    With more than one worker the loader is rebuilt around a ``DistributedSampler`` (shuffling iff
    the original sampler was random); ``loader.sampler.set_epoch`` stays available
    (``/root/reference/my_ray_module.py:151``). Batches are copied host->device through pinned memory
    on a side stream, ahead of the consumer.

    GPU workers, ``u8_fast_path``: a loader over a uint8 image dataset (FashionMNIST and friends, see
    ``_u8_image_source``) is replaced by ``data.U8BatchLoader`` — same batch sizes (tail batch included), pixels and labels
    kept as bytes, batches are ``data.U8Images`` that normalise themselves on the GPU (or are consumed raw by the
    engine-backed model of :func:`prepare_model`). ``transfer`` (default: ``$RTDC_INPUT_TRANSFER`` or ``"auto"``) picks how a
    shuffled batch reaches the device, see ``data.U8BatchLoader``: ``"resident"`` (dataset uploaded once, exact
    ``DistributedSampler`` shuffle by a gather kernel; what ``"auto"`` picks when the dataset fits ``$RTDC_RESIDENT_MAX_BYTES``,
    4 GiB by default), ``"stream"`` (dataset
    in pinned host memory, every step's pixels cross PCIe by DMA, two-level block shuffle) or ``"host_gather"`` (exact
    shuffle gathered by host threads).
    """
    from ...data.loader import DeviceLoader, U8BatchLoader, _claim_timed_loader, shard_data_loader
    import torch.utils.data as tud
    ctx = get_context()
    world, rank = ctx.get_world_size(), ctx.get_world_rank()
    device = get_device()
    timed = _claim_timed_loader()
    if u8_fast_path and move_to_device and device.type == "cuda":
        src = _u8_image_source(data_loader)
        if src is not None:
            imgs, labels, mean, std = src
            shard = add_dist_sampler and world > 1
            return U8BatchLoader(imgs, labels, data_loader.batch_size, device, shuffle=isinstance(data_loader.sampler, tud.RandomSampler),
                                 world=world if shard else 1, rank=rank if shard else 0, drop_last=data_loader.drop_last,
                                 mean=mean, std=std, wrap=True, timed=timed, lookahead=3,
                                 transfer=transfer or os.environ.get("RTDC_INPUT_TRANSFER", "auto"),
                                 resident_max_bytes=int(os.environ.get("RTDC_RESIDENT_MAX_BYTES", 4 << 30)))
    if add_dist_sampler and world > 1:
        data_loader = shard_data_loader(data_loader, world, rank)
    if move_to_device:
        return DeviceLoader(data_loader, device, prefetch=auto_transfer, timed=timed)
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
