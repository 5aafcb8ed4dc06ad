"""Worker-side session: ``report``, ``get_checkpoint``, ``get_context``.

Covers the calls the reference makes inside ``train_func_per_worker``:
``ray.train.report({...}, checkpoint=...)`` (``/root/reference/my_ray_module.py:203-205``) and
``ray.train.get_context().get_world_size()/get_world_rank()`` (``:149,177``), plus
``ray.train.get_checkpoint()`` which the north star requires for retry-resume.
"""
from __future__ import annotations

import os
import threading
import time
from typing import Any, Dict, Optional

from .checkpoint import Checkpoint
from .storage import StorageContext, checkpoint_dir_name


class TrainContext:
    def __init__(self, world_rank=0, world_size=1, local_rank=0, local_world_size=1, node_rank=0,
                 experiment_name="", trial_name="", trial_dir="", storage: Optional[StorageContext] = None,
                 use_gpu=False, metadata=None):
        self._world_rank, self._world_size = world_rank, world_size
        self._local_rank, self._local_world_size, self._node_rank = local_rank, local_world_size, node_rank
        self._experiment_name, self._trial_name, self._trial_dir = experiment_name, trial_name, trial_dir
        self._storage, self._use_gpu, self._metadata = storage, use_gpu, metadata or {}

    def get_world_size(self) -> int: return self._world_size
    def get_world_rank(self) -> int: return self._world_rank
    def get_local_rank(self) -> int: return self._local_rank
    def get_local_world_size(self) -> int: return self._local_world_size
    def get_node_rank(self) -> int: return self._node_rank
    def get_experiment_name(self) -> str: return self._experiment_name
    def get_trial_name(self) -> str: return self._trial_name
    def get_trial_id(self) -> str: return self._trial_name
    def get_trial_dir(self) -> str: return self._trial_dir
    def get_trial_resources(self): return {"GPU": 1} if self._use_gpu else {"CPU": 1}
    def get_metadata(self) -> Dict[str, Any]: return dict(self._metadata)
    def get_storage(self) -> Optional[StorageContext]: return self._storage


class _Session:
    """State of one worker inside a ``fit()``. ``link`` carries reports to whoever does the driver's
    bookkeeping (a socket to the driver process, or collectives in SPMD/torchrun mode)."""

    def __init__(self, context: TrainContext, link, start_checkpoint: Optional[Checkpoint], start_index: int = 0):
        self.context = context
        self.link = link
        self.latest_checkpoint = start_checkpoint
        self.report_index = start_index
        self.lock = threading.Lock()

    def report(self, metrics: Dict[str, Any], checkpoint=None, asynchronous: bool = False):
        if not isinstance(metrics, dict):
            raise TypeError("report(metrics=...) must be a dict")
        self.drain()                      # at most one report in flight; keeps indices ordered
        idx = self.report_index
        self.report_index += 1
        if asynchronous:
            def run():
                try:
                    self._do_report(idx, metrics, checkpoint)
                except BaseException as e:   # surfaced by the next report()/drain()
                    self._pending_error = e
            self._pending = threading.Thread(target=run, name="rtdc-report", daemon=True)
            self._pending.start()
            return
        self._do_report(idx, metrics, checkpoint)

    def drain(self):
        """Wait for an asynchronous report (checkpoint write + persist + driver barrier) to finish."""
        t = getattr(self, "_pending", None)
        if t is not None:
            t.join()
            self._pending = None
        err = getattr(self, "_pending_error", None)
        if err is not None:
            self._pending_error = None
            raise err

    def _do_report(self, idx: int, metrics: Dict[str, Any], checkpoint):
        ctx = self.context
        persisted_path = None
        t0 = time.time()
        if checkpoint is not None and hasattr(checkpoint, "result") and not isinstance(checkpoint, Checkpoint):
            checkpoint = checkpoint.result()          # ckpt.PendingCheckpoint: wait for the writer thread
        t1 = time.time()
        if checkpoint is not None:
            storage = ctx.get_storage()
            with checkpoint.as_directory() as src:
                persisted_path = storage.persist_checkpoint_files(src, idx, ctx.get_world_rank())
            metrics = dict(metrics)
            metrics.setdefault("_checkpoint_persist_s", round(time.time() - t1, 6))
        t2 = time.time()
        # blocks until every rank has reported this index and the bookkeeping is done (C5 barrier)
        self.link.report(rank=ctx.get_world_rank(), index=idx, metrics=metrics, checkpoint_path=persisted_path)
        t3 = time.time()
        self.phase_s = [a + b for a, b in zip(getattr(self, "phase_s", [0.0, 0.0, 0.0]), (t1 - t0, t2 - t1, t3 - t2))]
        if os.environ.get("RTDC_PROFILE_EPOCHS"):
            print(f"[rtdc] rank {ctx.get_world_rank()} report {idx}: wait for the checkpoint writer {(t1 - t0) * 1e3:.2f} ms, persist "
                  f"{(t2 - t1) * 1e3:.2f} ms, driver round trip {(t3 - t2) * 1e3:.2f} ms", flush=True)
        if self.link.any_checkpoint(idx):
            self.latest_checkpoint = Checkpoint.from_directory(ctx.get_storage().checkpoint_path(idx))
        _maybe_inject_fault(ctx.get_world_rank(), idx + 1)


_session: Optional[_Session] = None


def _set_session(s: Optional[_Session]):
    global _session
    _session = s


def _get_session(what: str) -> _Session:
    if _session is None:
        raise RuntimeError(f"`{what}` can only be called inside a train_loop_per_worker started by TorchTrainer.fit()")
    return _session


def in_session() -> bool:
    return _session is not None


def report(metrics: Dict[str, Any], *, checkpoint=None, checkpoint_upload_mode: str = "sync") -> None:
    """Report metrics (and optionally a checkpoint directory) for this step/epoch.

    Must be called by every worker the same number of times. The checkpoint's files are persisted to
    ``<storage_path>/<experiment>/<trial>/checkpoint_{index:06d}/`` and, in the default ``"sync"`` mode,
    the call returns once all workers have reported and retention (``CheckpointConfig.num_to_keep``)
    has been applied.

    ``checkpoint`` may be a ``Checkpoint`` or a ``ckpt.PendingCheckpoint`` (files still being written
    by the asynchronous checkpointer). With ``checkpoint_upload_mode="async"`` — pass it on every rank
    — waiting for the writer, persisting and the cross-worker barrier all happen on a background
    thread while training continues; the next ``report`` (or the end of the loop) waits for it.
    """
    if checkpoint_upload_mode not in ("sync", "async"):
        raise ValueError("checkpoint_upload_mode must be 'sync' or 'async'")
    _get_session("report").report(metrics, checkpoint, asynchronous=checkpoint_upload_mode == "async")


def get_checkpoint() -> Optional[Checkpoint]:
    """Latest checkpoint persisted by this run (``None`` on a fresh start).

    After a worker-group restart (``FailureConfig.max_failures``) or when the trainer was given
    ``resume_from_checkpoint``, this is where the training function finds its starting state.
    """
    s = _get_session("get_checkpoint")
    s.drain()
    return s.latest_checkpoint


def get_context() -> TrainContext:
    if _session is None:
        # outside fit(): single-process defaults, so library code can call it unconditionally
        return TrainContext()
    return _session.context


# ------------------------------------------------------------------------------------------------
# fault injection for tests: RTDC_FAULT="rank:report_index[:attempt]" kills that
# rank right after it completed the given number of reports (on the given attempt, default 0).
def _maybe_inject_fault(rank: int, reports_done: int):
    spec = os.environ.get("RTDC_FAULT")
    if not spec:
        return
    parts = spec.split(":")
    try:
        f_rank, f_idx = int(parts[0]), int(parts[1])
        f_attempt = int(parts[2]) if len(parts) > 2 else 0
    except (ValueError, IndexError):
        return
    attempt = int(os.environ.get("RTDC_ATTEMPT", "0"))
    if rank == f_rank and reports_done == f_idx and attempt == f_attempt:
        import sys
        print(f"[rtdc] injected fault: rank {rank} dies after report {reports_done} (attempt {attempt})", flush=True)
        sys.stdout.flush()
        os._exit(17)
