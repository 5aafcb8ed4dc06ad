"""Run / scaling / checkpoint / failure configuration objects.

Same names and fields as the Ray Train objects the reference constructs at
``/root/reference/my_ray_module.py:235-243`` (``RunConfig(checkpoint_config=CheckpointConfig(
num_to_keep=2), storage_path=..., verbose=1)``, ``ScalingConfig(num_workers, use_gpu)``).
"""
from __future__ import annotations

import dataclasses
import os
from typing import Any, Dict, Optional


@dataclasses.dataclass
class ScalingConfig:
    """How many worker processes to run and on what.

    ``use_gpu=True`` places one rank per GPU (``cuda:<local_rank>``) and selects the NVLink/NCCL
    communication stack; ``use_gpu=False`` runs CPU workers over gloo.
    """
    num_workers: int = 1
    use_gpu: bool = False
    resources_per_worker: Optional[Dict[str, float]] = None
    trainer_resources: Optional[Dict[str, float]] = None
    placement_strategy: str = "PACK"

    def __post_init__(self):
        if int(self.num_workers) < 1:
            raise ValueError(f"num_workers must be >= 1 (got {self.num_workers})")
        self.num_workers = int(self.num_workers)

    @property
    def num_gpus_per_worker(self) -> int:
        if not self.use_gpu:
            return 0
        if self.resources_per_worker and "GPU" in self.resources_per_worker:
            return int(self.resources_per_worker["GPU"])
        return 1

    @property
    def backend(self) -> str:
        return "nccl" if self.use_gpu else "gloo"


@dataclasses.dataclass
class CheckpointConfig:
    """Retention policy for persisted checkpoints.

    ``num_to_keep=None`` keeps everything. With no ``checkpoint_score_attribute`` the newest
    ``num_to_keep`` are kept (the reference uses ``num_to_keep=2``); with a score attribute the best
    ones by that metric are kept, and the latest checkpoint is never deleted.
    """
    num_to_keep: Optional[int] = None
    checkpoint_score_attribute: Optional[str] = None
    checkpoint_score_order: str = "max"

    def __post_init__(self):
        if self.num_to_keep is not None and self.num_to_keep <= 0:
            raise ValueError(f"num_to_keep must be None or a positive integer (got {self.num_to_keep})")
        if self.checkpoint_score_order not in ("max", "min"):
            raise ValueError("checkpoint_score_order must be 'max' or 'min'")


@dataclasses.dataclass
class FailureConfig:
    """``max_failures`` worker-group restarts before ``fit()`` gives up (-1 = unlimited). A restarted
    group sees the latest persisted checkpoint through ``train.get_checkpoint()``."""
    max_failures: int = 0
    fail_fast: bool = False


@dataclasses.dataclass
class RunConfig:
    name: Optional[str] = None
    storage_path: Optional[str] = None
    storage_filesystem: Optional[Any] = None     # a pyarrow.fs.FileSystem; storage_path is then a path inside it
    checkpoint_config: Optional[CheckpointConfig] = None
    failure_config: Optional[FailureConfig] = None
    verbose: int = 1
    log_to_file: bool = False
    worker_start_timeout_s: float = 300.0   # ref: @metaflow_ray(all_nodes_started_timeout=300), train_flow.py:42
    # fit() fails the attempt when no worker message (report, heartbeat) arrives for this long: a rank stuck in host code
    # would otherwise stall the group forever (device-side waits are bounded separately). Workers send a heartbeat every
    # worker_heartbeat_s from a background thread, so a long epoch without report() calls does not trip it. None disables.
    worker_report_timeout_s: Optional[float] = 600.0
    worker_heartbeat_s: float = 5.0
    callbacks: Optional[list] = None

    def __post_init__(self):
        if self.checkpoint_config is None:
            self.checkpoint_config = CheckpointConfig()
        if self.failure_config is None:
            self.failure_config = FailureConfig()
        if self.storage_path is None:
            self.storage_path = os.environ.get("RTDC_STORAGE_PATH", os.path.expanduser("~/rtdc_results"))
        self.storage_path = str(self.storage_path)


def asdict_shallow(obj: Any) -> Dict[str, Any]:
    return {f.name: getattr(obj, f.name) for f in dataclasses.fields(obj)}
