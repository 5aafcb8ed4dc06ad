"""Run-directory layout, checkpoint persistence and retention.

Layout (what ``--from-run`` and the eval flow consume)::

    <storage_path>/<experiment_name>/<trial_name>/
        checkpoint_000000/ latest_model.pt [best_model.pt]
        checkpoint_000001/ ...
        result.json  progress.csv  params.json

Reference behaviour being covered: ``ray.train.report(metrics, checkpoint=...)`` copying the worker's
directory into ``checkpoint_{idx:06d}`` and ``CheckpointConfig(num_to_keep=2)`` pruning
(``/root/reference/my_ray_module.py:202-205,236``). Unlike the reference, same-named files from
different ranks do not race: rank 0's copy always wins, other ranks only fill in files rank 0 did not
provide, and every file appears atomically.
"""
from __future__ import annotations

import csv
import datetime as _dt
import json
import os
import shutil
import threading
import uuid
from typing import Any, Dict, List, Optional, Tuple

from .checkpoint import Checkpoint, _strip_scheme
from .config import CheckpointConfig


def _timestamp() -> str:
    return _dt.datetime.now().strftime("%Y-%m-%d_%H-%M-%S")


def checkpoint_dir_name(index: int) -> str:
    return f"checkpoint_{index:06d}"


class StorageContext:
    """Resolves where a run lives and moves checkpoint files there.

    ``storage_path`` may be a local path / ``file://`` URI (the run directory is built in place), or anything ``pyarrow.fs``
    can address — an ``s3://`` / ``gs://`` / ``hdfs://`` URI, or any path together with an explicit ``storage_filesystem``
    (``RunConfig(storage_filesystem=...)``, as in Ray). Remote runs are staged in a local directory that all workers of the
    node share, and the driver mirrors every completed report (checkpoint directory, result.json, progress.csv) to the remote
    filesystem and applies retention there too; ``Result.path`` and the ``Checkpoint`` handles then point at the remote copy
    (ref: ``storage_path=current.ray_storage_path``, a datastore URI, train_flow.py:65 / README.md:13-15)."""

    def __init__(self, storage_path: str, experiment_name: Optional[str] = None, trial_name: Optional[str] = None,
                 storage_filesystem: Optional[Any] = None):
        local, scheme = _strip_scheme(storage_path)
        self.remote_fs, self.remote_root = None, None
        if storage_filesystem is not None or scheme not in ("local", "file"):
            import hashlib
            import pyarrow.fs as pafs
            if storage_filesystem is not None:
                self.remote_fs, self.remote_root = storage_filesystem, str(storage_path).rstrip("/")
            else:
                try:
                    self.remote_fs, self.remote_root = pafs.FileSystem.from_uri(storage_path)
                except Exception as e:
                    raise ValueError(f"storage_path {storage_path!r} is neither a local path nor a URI pyarrow.fs can open: {e}") from e
                self.remote_root = self.remote_root.rstrip("/")
            tag = hashlib.sha1(f"{type(self.remote_fs).__name__}:{storage_path}".encode()).hexdigest()[:12]
            local = os.path.join(os.environ.get("RTDC_STAGING_DIR", os.path.expanduser("~/.cache/rtdc_staging")), tag)
        self.storage_uri = str(storage_path)
        self.storage_path = os.path.abspath(os.path.expanduser(local))
        self.experiment_name = experiment_name or f"TorchTrainer_{_timestamp()}"
        self.trial_name = trial_name or f"TorchTrainer_{uuid.uuid4().hex[:5]}_00000_0_{_timestamp()}"

    # -------------------------------------------------------------- remote mirror (driver side)
    @property
    def is_remote(self) -> bool:
        return self.remote_fs is not None

    @property
    def remote_trial_path(self) -> Optional[str]:
        return f"{self.remote_root}/{self.experiment_name}/{self.trial_name}" if self.is_remote else None

    @property
    def result_path(self) -> str:
        """What ``Result.path`` reports: the remote trial directory for remote storage, else the local one."""
        if not self.is_remote:
            return self.trial_path
        return self.storage_uri.rstrip("/") + f"/{self.experiment_name}/{self.trial_name}"

    def mirror_checkpoint(self, index: int) -> "Checkpoint":
        """Upload ``checkpoint_{index}`` of the staging directory; return the handle of the remote copy."""
        import pyarrow.fs as pafs
        dst = f"{self.remote_trial_path}/{checkpoint_dir_name(index)}"
        self.remote_fs.create_dir(dst, recursive=True)
        pafs.copy_files(self.checkpoint_path(index), dst, source_filesystem=pafs.LocalFileSystem(), destination_filesystem=self.remote_fs)
        return Checkpoint(dst, filesystem=self.remote_fs)

    def mirror_logs(self):
        import pyarrow.fs as pafs
        self.remote_fs.create_dir(self.remote_trial_path, recursive=True)
        for name in ("result.json", "progress.csv", "params.json"):
            src = os.path.join(self.trial_path, name)
            if os.path.exists(src):
                pafs.copy_files(src, f"{self.remote_trial_path}/{name}", source_filesystem=pafs.LocalFileSystem(),
                                destination_filesystem=self.remote_fs)

    def delete_checkpoint(self, checkpoint: "Checkpoint"):
        """Retention: remove a checkpoint directory (and, for remote storage, its staged local twin)."""
        local = checkpoint._local_path()
        if local and os.path.isdir(local):
            shutil.rmtree(local, ignore_errors=True)
            return
        if self.is_remote:
            try:
                self.remote_fs.delete_dir(checkpoint.path)
            except Exception:
                pass
            twin = os.path.join(self.trial_path, os.path.basename(checkpoint.path.rstrip("/")))
            shutil.rmtree(twin, ignore_errors=True)

    @property
    def experiment_path(self) -> str:
        return os.path.join(self.storage_path, self.experiment_name)

    @property
    def trial_path(self) -> str:
        return os.path.join(self.experiment_path, self.trial_name)

    def checkpoint_path(self, index: int) -> str:
        return os.path.join(self.trial_path, checkpoint_dir_name(index))

    def make_dirs(self):
        os.makedirs(self.trial_path, exist_ok=True)

    # -------------------------------------------------------------- worker side
    def persist_checkpoint_files(self, src_dir: str, index: int, rank: int) -> str:
        """Copy every file under ``src_dir`` into the run's ``checkpoint_{index}`` directory.

        Each file is written to a temp name and then published atomically. Rank 0 replaces whatever is
        there; other ranks publish with ``link`` (fails if the name exists), so for files several ranks
        provide the result is always rank 0's.
        """
        dst_dir = self.checkpoint_path(index)
        os.makedirs(dst_dir, exist_ok=True)
        src_dir = os.path.abspath(src_dir)
        if src_dir == os.path.abspath(dst_dir):
            return dst_dir
        for root, _dirs, files in os.walk(src_dir):
            rel = os.path.relpath(root, src_dir)
            out_root = dst_dir if rel == "." else os.path.join(dst_dir, rel)
            os.makedirs(out_root, exist_ok=True)
            for name in files:
                src = os.path.join(root, name)
                dst = os.path.join(out_root, name)
                if rank != 0 and os.path.exists(dst):
                    continue
                tmp = os.path.join(out_root, f".{name}.rank{rank}.{os.getpid()}.tmp")
                try:
                    os.link(src, tmp)                 # same filesystem: publishing a checkpoint file costs no copy
                except OSError:
                    shutil.copyfile(src, tmp)
                if rank == 0:
                    os.replace(tmp, dst)
                else:
                    try:
                        os.link(tmp, dst)
                    except FileExistsError:
                        pass
                    finally:
                        os.unlink(tmp)
        return dst_dir


class CheckpointManager:
    """Driver-side registry of persisted checkpoints + retention (``num_to_keep``)."""

    def __init__(self, config: CheckpointConfig, delete=None):
        self.config = config
        self._delete = delete            # how to remove a checkpoint that fell out of retention (StorageContext.delete_checkpoint)
        self._entries: List[Tuple[Checkpoint, Dict[str, Any], int]] = []   # (checkpoint, metrics, index)
        self._lock = threading.Lock()

    @property
    def latest(self) -> Optional[Checkpoint]:
        return self._entries[-1][0] if self._entries else None

    @property
    def best_checkpoints(self) -> List[Tuple[Checkpoint, Dict[str, Any]]]:
        return [(c, m) for c, m, _ in self._entries]

    def _score(self, entry):
        attr = self.config.checkpoint_score_attribute
        _c, metrics, index = entry
        if attr is None:
            return (1, index)
        if attr not in metrics:
            return (0, index)   # unscored checkpoints go first
        v = metrics[attr]
        return (1, v if self.config.checkpoint_score_order == "max" else -v)

    def register(self, checkpoint: Checkpoint, metrics: Dict[str, Any], index: int) -> List[Checkpoint]:
        """Record a new checkpoint; delete and return the ones that fall out of retention."""
        with self._lock:
            self._entries.append((checkpoint, dict(metrics), index))
            k = self.config.num_to_keep
            if k is None or len(self._entries) <= k:
                return []
            latest = self._entries[-1]
            ranked = sorted(self._entries, key=self._score)          # worst first
            doomed = []
            for e in ranked:
                if len(self._entries) - len(doomed) <= k:
                    break
                if e is latest:
                    continue                                          # never delete the newest
                doomed.append(e)
            for e in doomed:
                self._entries.remove(e)
                if self._delete is not None:
                    self._delete(e[0])
                else:
                    local = e[0]._local_path()
                    if local and os.path.isdir(local):
                        shutil.rmtree(local, ignore_errors=True)
            return [e[0] for e in doomed]


class ResultLogger:
    """Appends one row per report to ``result.json`` (JSON lines) and ``progress.csv``."""

    def __init__(self, trial_path: str, params: Optional[Dict[str, Any]] = None):
        self.trial_path = trial_path
        self._csv_fields: Optional[List[str]] = None
        os.makedirs(trial_path, exist_ok=True)
        if params is not None:
            with open(os.path.join(trial_path, "params.json"), "w") as f:
                json.dump(params, f, default=repr, indent=1)

    def log(self, row: Dict[str, Any]):
        flat = {k: v for k, v in row.items() if isinstance(v, (int, float, str, bool)) or v is None}
        with open(os.path.join(self.trial_path, "result.json"), "a") as f:
            f.write(json.dumps(flat) + "\n")
        path = os.path.join(self.trial_path, "progress.csv")
        if self._csv_fields is None:
            self._csv_fields = list(flat.keys())
            with open(path, "w", newline="") as f:
                csv.DictWriter(f, fieldnames=self._csv_fields).writeheader()
        with open(path, "a", newline="") as f:
            csv.DictWriter(f, fieldnames=self._csv_fields, extrasaction="ignore").writerow(flat)
