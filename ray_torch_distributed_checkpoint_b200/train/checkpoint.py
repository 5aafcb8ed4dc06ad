"""``Checkpoint``: a picklable handle to a directory of checkpoint files.

Mirrors what the reference uses from ``ray.train.Checkpoint``: ``Checkpoint.from_directory(dir)``
(``/root/reference/my_ray_module.py:202``), ``.path`` (``:133``), ``with ckpt.as_directory() as d``
(``:254``); it rides inside the pickled ``Result`` artifact between flow runs
(``/root/reference/train_flow.py:70-77``).
"""
from __future__ import annotations

import contextlib
import json
import os
import shutil
import tempfile
import uuid
from typing import Any, Dict, Iterator, Optional

_METADATA_FILE = ".metadata.json"


def _strip_scheme(path: str):
    if path.startswith("file://"):
        return path[len("file://"):], "file"
    if "://" in path:
        scheme = path.split("://", 1)[0]
        return path, scheme
    return path, "local"


class Checkpoint:
    """Reference to a checkpoint directory on a filesystem reachable from this process.

    Local paths and ``file://`` URIs are supported natively; other schemes go through
    ``pyarrow.fs`` when it can resolve them.
    """

    def __init__(self, path: str, filesystem: Optional[Any] = None):
        self.path = str(path)
        self.filesystem = filesystem
        self._uuid = uuid.uuid4().hex

    # ------------------------------------------------------------------ construction
    @classmethod
    def from_directory(cls, path: str) -> "Checkpoint":
        return cls(os.path.abspath(os.fspath(path)))

    # ------------------------------------------------------------------ access
    def _local_path(self) -> Optional[str]:
        p, scheme = _strip_scheme(self.path)
        return p if scheme in ("local", "file") and self.filesystem is None else None

    def _download(self, dst: str) -> str:
        local = self._local_path()
        os.makedirs(dst, exist_ok=True)
        if local is not None:
            if os.path.abspath(local) != os.path.abspath(dst):
                shutil.copytree(local, dst, dirs_exist_ok=True)
            return dst
        import pyarrow.fs as pafs
        fs = self.filesystem
        path = self.path
        if fs is None:
            fs, path = pafs.FileSystem.from_uri(self.path)
        pafs.copy_files(path, dst, source_filesystem=fs, destination_filesystem=pafs.LocalFileSystem())
        return dst

    def to_directory(self, path: Optional[str] = None) -> str:
        """Materialise the checkpoint files in ``path`` (a fresh temp dir by default)."""
        if path is None:
            path = os.path.join(tempfile.gettempdir(), f"rtdc_ckpt_{self._uuid}_{os.getpid()}")
        return self._download(path)

    @contextlib.contextmanager
    def as_directory(self) -> Iterator[str]:
        """Yield a local directory with the checkpoint files. A local checkpoint is yielded in place
        (no copy); a remote one is downloaded to a temp dir that is removed afterwards."""
        local = self._local_path()
        if local is not None:
            yield local
            return
        tmp = tempfile.mkdtemp(prefix="rtdc_ckpt_")
        try:
            yield self._download(tmp)
        finally:
            shutil.rmtree(tmp, ignore_errors=True)

    # ------------------------------------------------------------------ metadata
    def _metadata_path(self) -> str:
        local = self._local_path()
        if local is None:
            raise NotImplementedError("metadata is only supported for local checkpoints")
        return os.path.join(local, _METADATA_FILE)

    def get_metadata(self) -> Dict[str, Any]:
        p = self._metadata_path()
        if not os.path.exists(p):
            return {}
        with open(p) as f:
            return json.load(f)

    def set_metadata(self, metadata: Dict[str, Any]) -> None:
        p = self._metadata_path()
        tmp = p + f".tmp{os.getpid()}"
        with open(tmp, "w") as f:
            json.dump(metadata, f)
        os.replace(tmp, p)

    def update_metadata(self, metadata: Dict[str, Any]) -> None:
        cur = self.get_metadata()
        cur.update(metadata)
        self.set_metadata(cur)

    # ------------------------------------------------------------------ dunder
    def __repr__(self) -> str:
        return f"Checkpoint(path={self.path})" if self.filesystem is None else f"Checkpoint(filesystem={type(self.filesystem).__name__}, path={self.path})"

    def __eq__(self, other) -> bool:
        return isinstance(other, Checkpoint) and other.path == self.path

    def __hash__(self) -> int:
        return hash(self.path)

    def __getstate__(self):
        return {"path": self.path, "filesystem": self.filesystem, "_uuid": self._uuid}

    def __setstate__(self, state):
        self.path = state["path"]
        self.filesystem = state.get("filesystem")
        self._uuid = state.get("_uuid", uuid.uuid4().hex)
