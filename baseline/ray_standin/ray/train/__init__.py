"""ray.train stand-in on stock torch.distributed (SPMD: every rank runs the driver code)."""
from __future__ import annotations

import contextlib
import os
import shutil
import time

import torch
import torch.distributed as dist


class ScalingConfig:
    def __init__(self, num_workers=1, use_gpu=False, **_):
        self.num_workers, self.use_gpu = num_workers, use_gpu


class CheckpointConfig:
    def __init__(self, num_to_keep=None, **_):
        self.num_to_keep = num_to_keep


class RunConfig:
    def __init__(self, checkpoint_config=None, storage_path=None, verbose=1, name=None, **_):
        self.checkpoint_config = checkpoint_config or CheckpointConfig()
        self.storage_path = storage_path or os.path.expanduser("~/ray_results")
        self.verbose, self.name = verbose, name


class Checkpoint:
    def __init__(self, path):
        self.path = path

    @classmethod
    def from_directory(cls, path):
        return cls(os.path.abspath(path))

    @contextlib.contextmanager
    def as_directory(self):
        yield self.path


class Result:
    def __init__(self, metrics, checkpoint, path):
        self.metrics, self.checkpoint, self.path, self.error = metrics, checkpoint, path, None

    def __repr__(self):
        return f"Result(metrics={self.metrics}, path={self.path!r}, checkpoint={self.checkpoint and self.checkpoint.path})"


class _Context:
    def get_world_size(self):
        return dist.get_world_size() if dist.is_initialized() else 1

    def get_world_rank(self):
        return dist.get_rank() if dist.is_initialized() else 0

    def get_local_rank(self):
        return int(os.environ.get("LOCAL_RANK", 0))


_state = {"run_dir": None, "index": 0, "kept": [], "num_to_keep": None, "last_metrics": None, "last_ckpt": None,
          "use_gpu": False}


def get_context():
    return _Context()


def get_checkpoint():
    return _state["last_ckpt"]


def report(metrics, checkpoint=None):
    """Every rank copies its checkpoint files into the shared checkpoint_{idx:06d} dir, then all ranks meet."""
    idx = _state["index"]
    _state["index"] += 1
    dst = os.path.join(_state["run_dir"], f"checkpoint_{idx:06d}")
    if checkpoint is not None:
        os.makedirs(dst, exist_ok=True)
        for name in os.listdir(checkpoint.path):
            shutil.copy2(os.path.join(checkpoint.path, name), os.path.join(dst, name))
    if dist.is_initialized():
        dist.barrier()
    if checkpoint is not None:
        _state["last_ckpt"] = Checkpoint(dst)
        _state["kept"].append(dst)
        k = _state["num_to_keep"]
        rank = dist.get_rank() if dist.is_initialized() else 0
        while k is not None and len(_state["kept"]) > k:
            old = _state["kept"].pop(0)
            if rank == 0:
                shutil.rmtree(old, ignore_errors=True)
    _state["last_metrics"] = dict(metrics)


from . import torch as torch  # noqa: E402,F401  (ray.train.torch)
