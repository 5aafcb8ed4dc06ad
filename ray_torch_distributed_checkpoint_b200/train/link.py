"""Control plane between the driver (``TorchTrainer.fit``) and its workers.

* ``DriverServer`` / ``SocketLink``: an AF_UNIX ``multiprocessing.connection`` channel per worker.
  ``report()`` on a worker returns only after the driver has seen that report index from every rank
  and finished its bookkeeping — the driver-mediated barrier of ``ray.train.report``.
* ``SpmdLink``: the same contract when the user launched all ranks themselves (torchrun): rank 0 does
  the bookkeeping, the ranks meet in ``torch.distributed`` collectives.
"""
from __future__ import annotations

import os
import queue
import threading
from multiprocessing.connection import Client, Listener
from typing import Any, Callable, Dict, List, Optional


class WorkerFailure(Exception):
    def __init__(self, rank: int, message: str):
        super().__init__(f"worker rank {rank} failed:\n{message}")
        self.rank = rank
        self.message = message


class DriverServer:
    """Accepts one connection per rank and turns their messages into an event queue."""

    def __init__(self, socket_path: str, world_size: int, authkey: bytes):
        self.world_size = world_size
        self.listener = Listener(address=socket_path, family="AF_UNIX", authkey=authkey)
        self.conns: Dict[int, Any] = {}
        self.events: "queue.Queue[dict]" = queue.Queue()
        self._closed = False
        self._accept_thread = threading.Thread(target=self._accept_loop, daemon=True)
        self._accept_thread.start()

    def _accept_loop(self):
        while not self._closed and len(self.conns) < self.world_size:
            try:
                conn = self.listener.accept()
            except (OSError, EOFError):
                return
            try:
                hello = conn.recv()
            except (EOFError, OSError):
                continue
            rank = hello["rank"]
            self.conns[rank] = conn
            self.events.put({"type": "hello", "rank": rank, "pid": hello.get("pid")})
            threading.Thread(target=self._reader, args=(rank, conn), daemon=True).start()

    def _reader(self, rank: int, conn):
        while not self._closed:
            try:
                msg = conn.recv()
            except (EOFError, OSError):
                self.events.put({"type": "eof", "rank": rank})
                return
            msg["rank"] = rank
            self.events.put(msg)
            if msg.get("type") in ("done", "error"):
                return

    def broadcast(self, msg: dict):
        for conn in list(self.conns.values()):
            try:
                conn.send(msg)
            except (OSError, EOFError, BrokenPipeError):
                pass

    def close(self):
        self._closed = True
        try:
            self.listener.close()
        except OSError:
            pass
        for conn in list(self.conns.values()):
            try:
                conn.close()
            except OSError:
                pass


class SocketLink:
    """Worker end of the channel."""

    def __init__(self, socket_path: str, authkey: bytes, rank: int, heartbeat_s: float = 0.0):
        self.rank = rank
        self.conn = Client(address=socket_path, family="AF_UNIX", authkey=authkey)
        self._send_lock = threading.Lock()       # the heartbeat thread shares the connection's send side
        self._send({"type": "hello", "rank": rank, "pid": os.getpid()})
        self._ckpt_flags: Dict[int, bool] = {}
        self._stop = threading.Event()
        if heartbeat_s and heartbeat_s > 0:
            threading.Thread(target=self._beat, args=(float(heartbeat_s),), name="rtdc-heartbeat", daemon=True).start()

    def _send(self, msg: dict):
        with self._send_lock:
            self.conn.send(msg)

    def _beat(self, period: float):
        """Tell the driver this process is alive (a host-side hang stops the beats: a thread stuck holding the GIL
        forever, a dead process; the driver then fails the attempt after RunConfig.worker_report_timeout_s)."""
        while not self._stop.wait(period):
            try:
                self._send({"type": "heartbeat"})
            except (OSError, EOFError, BrokenPipeError, ValueError):
                return

    def report(self, rank: int, index: int, metrics: Dict[str, Any], checkpoint_path: Optional[str]):
        self._send({"type": "report", "index": index, "metrics": metrics, "checkpoint_path": checkpoint_path})
        while True:
            msg = self.conn.recv()          # blocks: this is the barrier
            if msg["type"] == "ack" and msg["index"] == index:
                self._ckpt_flags[index] = bool(msg.get("has_checkpoint"))
                return
            if msg["type"] == "abort":
                raise SystemExit(3)

    def any_checkpoint(self, index: int) -> bool:
        return self._ckpt_flags.pop(index, False)

    def done(self):
        self._stop.set()
        self._send({"type": "done"})

    def error(self, tb: str):
        self._stop.set()
        try:
            self._send({"type": "error", "traceback": tb})
        except (OSError, EOFError, BrokenPipeError):
            pass

    def close(self):
        self._stop.set()
        try:
            self.conn.close()
        except OSError:
            pass


class SpmdLink:
    """All ranks were launched by someone else (torchrun); bookkeeping runs on rank 0 in-process."""

    def __init__(self, rank: int, world_size: int, on_complete: Optional[Callable[[int, List[dict]], bool]]):
        self.rank, self.world_size = rank, world_size
        self.on_complete = on_complete       # rank 0 only: (index, per-rank reports) -> has_checkpoint
        self._ckpt_flags: Dict[int, bool] = {}
        # Reports may run on a background thread (checkpoint_upload_mode="async") while the main thread issues gradient
        # collectives on the default group: collectives from two threads on ONE group can be ordered differently on
        # different ranks. The report traffic therefore gets a gloo group of its own (collective call: every rank
        # constructs the link at the same point of fit()).
        self.group = None
        if world_size > 1:
            import torch.distributed as dist
            self.group = dist.new_group(backend="gloo")

    def report(self, rank: int, index: int, metrics: Dict[str, Any], checkpoint_path: Optional[str]):
        import torch.distributed as dist
        entry = {"rank": rank, "index": index, "metrics": metrics, "checkpoint_path": checkpoint_path}
        if self.world_size > 1:
            gathered: List[Optional[dict]] = [None] * self.world_size if rank == 0 else None
            dist.gather_object(entry, gathered, dst=0, group=self.group)
        else:
            gathered = [entry]
        flag = [False]
        if rank == 0:
            flag[0] = bool(self.on_complete(index, gathered))
        if self.world_size > 1:
            dist.broadcast_object_list(flag, src=0, group=self.group)     # doubles as the release barrier
        self._ckpt_flags[index] = flag[0]

    def any_checkpoint(self, index: int) -> bool:
        return self._ckpt_flags.pop(index, False)

    def done(self): pass
    def error(self, tb: str): pass

    def close(self):
        if self.group is not None:
            import torch.distributed as dist
            try:
                dist.destroy_process_group(self.group)
            except Exception:
                pass
            self.group = None
