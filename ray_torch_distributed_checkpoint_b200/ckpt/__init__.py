"""Checkpoint writing: asynchronous device snapshots, atomic ``torch.save``-compatible files, and
the reference's file names (``/root/reference/my_ray_module.py:27-28``)."""
from .async_writer import AsyncCheckpointer, PendingCheckpoint, atomic_torch_save, wait_all

BEST_CHECKPOINT_FILENAME = "best_model.pt"
LATEST_CHECKPOINT_FILENAME = "latest_model.pt"

__all__ = ["AsyncCheckpointer", "PendingCheckpoint", "atomic_torch_save", "wait_all",
           "BEST_CHECKPOINT_FILENAME", "LATEST_CHECKPOINT_FILENAME"]
