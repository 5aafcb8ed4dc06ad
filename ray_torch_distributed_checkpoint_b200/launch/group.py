"""Spawn and supervise the worker processes of one ``fit()`` attempt.

Reference behaviour covered: what ``TorchTrainer.fit()`` does implicitly at
``/root/reference/my_ray_module.py:244-250`` (start ``num_workers`` workers, give them
``RANK/LOCAL_RANK/WORLD_SIZE/MASTER_ADDR/MASTER_PORT``) and the gang semantics of
``@metaflow_ray`` + ``num_parallel`` (``/root/reference/train_flow.py:39,42``).
"""
from __future__ import annotations

import os
import signal
import socket
import subprocess
import sys
import time
from pathlib import Path
from typing import Dict, List, Optional

_LAUNCHER = Path(__file__).resolve().parent.parent / "lib" / "rtdc-launch"


def find_free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def native_launcher_path() -> Optional[str]:
    if os.environ.get("RTDC_NO_NATIVE_LAUNCHER"):
        return None
    return str(_LAUNCHER) if _LAUNCHER.exists() and os.access(_LAUNCHER, os.X_OK) else None


class WorkerGroup:
    def __init__(self, num_workers: int, argv: List[str], env: Dict[str, str], master_port: Optional[int] = None,
                 bind_cores: bool = True):
        self.num_workers = num_workers
        self.argv = list(argv)
        self.env = dict(env)
        self.master_port = master_port or find_free_port()
        self.bind_cores = bind_cores
        self.procs: List[subprocess.Popen] = []
        self.native = native_launcher_path() is not None

    def start(self):
        base_env = dict(os.environ)
        base_env.update(self.env)
        base_env["RTDC_WORKER"] = "1"
        launcher = native_launcher_path()
        if launcher:
            cmd = [launcher, "--nproc", str(self.num_workers), "--master-addr", "127.0.0.1",
                   "--master-port", str(self.master_port)]
            if self.bind_cores:
                cmd.append("--bind-cores")
            cmd += ["--"] + self.argv
            self.procs = [subprocess.Popen(cmd, env=base_env, start_new_session=True)]
            return
        for rank in range(self.num_workers):
            env = dict(base_env)
            env.update({"RANK": str(rank), "LOCAL_RANK": str(rank), "WORLD_SIZE": str(self.num_workers),
                        "LOCAL_WORLD_SIZE": str(self.num_workers), "MASTER_ADDR": "127.0.0.1",
                        "MASTER_PORT": str(self.master_port), "NODE_RANK": "0"})
            self.procs.append(subprocess.Popen(self.argv, env=env, start_new_session=True))

    def poll(self) -> Optional[int]:
        """None while everything runs (or finished with 0 and others still run); the first non-zero
        exit code as soon as any process failed; 0 when all exited cleanly."""
        codes = [p.poll() for p in self.procs]
        for c in codes:
            if c is not None and c != 0:
                return c
        if all(c == 0 for c in codes):
            return 0
        return None

    def shutdown(self, grace_s: float = 3.0):
        for p in self.procs:
            if p.poll() is None:
                try:
                    os.killpg(p.pid, signal.SIGTERM)
                except (ProcessLookupError, PermissionError):
                    pass
        deadline = time.time() + grace_s
        for p in self.procs:
            try:
                p.wait(timeout=max(0.0, deadline - time.time()))
            except subprocess.TimeoutExpired:
                try:
                    os.killpg(p.pid, signal.SIGKILL)
                except (ProcessLookupError, PermissionError):
                    pass
                p.wait()

    def wait(self, timeout: Optional[float] = None) -> int:
        deadline = None if timeout is None else time.time() + timeout
        while True:
            c = self.poll()
            if c is not None:
                return c
            if deadline is not None and time.time() > deadline:
                raise TimeoutError("worker group did not exit in time")
            time.sleep(0.02)


def worker_argv(payload_path: str) -> List[str]:
    return [sys.executable, "-m", "ray_torch_distributed_checkpoint_b200.launch.worker", payload_path]
