"""``TorchTrainer``: run a per-worker training function on N ranks and collect a ``Result``.

Same call shape as the reference's use of Ray Train
(``/root/reference/my_ray_module.py:244-250``)::

    trainer = TorchTrainer(train_loop_per_worker=fn, train_loop_config=cfg,
                           scaling_config=ScalingConfig(num_workers=N, use_gpu=True),
                           run_config=RunConfig(storage_path=..., checkpoint_config=CheckpointConfig(num_to_keep=2)))
    result = trainer.fit()

Two execution modes:

* **driver mode** (default): ``fit()`` spawns ``num_workers`` processes on this node through the
  native launcher (one per GPU when ``use_gpu``), gives them the rendezvous environment, and serves
  their ``report()`` calls over a local socket.
* **SPMD mode**: when the calling process is already one rank of a torchrun-style launch
  (``RANK``/``WORLD_SIZE`` set and not started by us), every rank calls ``fit()``, the loop runs
  in-process and rank 0 does the bookkeeping.
"""
from __future__ import annotations

import inspect
import os
import shutil
import sys
import tempfile
import time
import traceback
from typing import Any, Callable, Dict, List, Optional

from .checkpoint import Checkpoint
from .config import CheckpointConfig, FailureConfig, RunConfig, ScalingConfig, asdict_shallow
from .link import DriverServer, SpmdLink
from .result import Result
from .session import TrainContext, _Session, _set_session
from .storage import CheckpointManager, ResultLogger, StorageContext, checkpoint_dir_name


class TrainingFailedError(RuntimeError):
    """Raised by ``fit()`` when the worker group failed more often than ``FailureConfig.max_failures``."""


def _log(verbose: int, msg: str):
    if verbose:
        print(f"[rtdc] {msg}", flush=True)


def _call_train_fn(fn: Callable, config: Optional[Dict[str, Any]]):
    try:
        nparams = len([p for p in inspect.signature(fn).parameters.values()
                       if p.default is inspect.Parameter.empty and p.kind in (p.POSITIONAL_ONLY, p.POSITIONAL_OR_KEYWORD)])
    except (TypeError, ValueError):
        nparams = 1
    return fn(config if config is not None else {}) if nparams >= 1 else fn()


class _Bookkeeper:
    """What the driver does each time all ranks have reported an index."""

    def __init__(self, storage: StorageContext, run_config: RunConfig, params: Dict[str, Any]):
        self.storage = storage
        self.manager = CheckpointManager(run_config.checkpoint_config, delete=storage.delete_checkpoint)
        self.logger = ResultLogger(storage.trial_path, params)
        self.history: List[Dict[str, Any]] = []
        self.start_time = time.time()
        self.verbose = run_config.verbose
        self.callbacks = run_config.callbacks or []

    def complete(self, index: int, reports: List[dict]) -> bool:
        reports = sorted(reports, key=lambda r: r["rank"])
        metrics = dict(reports[0]["metrics"])          # rank 0 is authoritative
        has_ckpt = any(r.get("checkpoint_path") for r in reports)
        row = dict(metrics)
        row.update({"training_iteration": index + 1, "time_total_s": round(time.time() - self.start_time, 3),
                    "timestamp": int(time.time())})
        if has_ckpt:
            row["checkpoint_dir_name"] = checkpoint_dir_name(index)
        self.history.append(row)
        self.logger.log(row)
        if has_ckpt:
            if self.storage.is_remote:
                ckpt = self.storage.mirror_checkpoint(index)
            else:
                ckpt = Checkpoint.from_directory(self.storage.checkpoint_path(index))
            removed = self.manager.register(ckpt, row, index)
            for c in removed:
                _log(self.verbose > 1, f"retention: removed {c.path}")
        shown = {k: v for k, v in metrics.items() if not k.startswith("_")}
        _log(self.verbose, f"report {index + 1}: {shown}" + (f"  checkpoint -> {checkpoint_dir_name(index)}" if has_ckpt else ""))
        if self.storage.is_remote:
            self.storage.mirror_logs()
        for cb in self.callbacks:
            cb(index, row)
        return has_ckpt

    def result(self, error: Optional[Exception] = None) -> Result:
        return Result(metrics=self.history[-1] if self.history else None, checkpoint=self.manager.latest,
                      error=error, path=self.storage.result_path, metrics_history=list(self.history),
                      best_checkpoints=self.manager.best_checkpoints)


class TorchTrainer:
    def __init__(self, train_loop_per_worker: Callable, *, train_loop_config: Optional[Dict[str, Any]] = None,
                 scaling_config: Optional[ScalingConfig] = None, run_config: Optional[RunConfig] = None,
                 resume_from_checkpoint: Optional[Checkpoint] = None, metadata: Optional[Dict[str, Any]] = None,
                 torch_config: Optional[Any] = None, datasets: Optional[Dict[str, Any]] = None):
        if not callable(train_loop_per_worker):
            raise TypeError("train_loop_per_worker must be callable")
        self.train_loop_per_worker = train_loop_per_worker
        self.train_loop_config = train_loop_config
        self.scaling_config = scaling_config or ScalingConfig()
        self.run_config = run_config or RunConfig()
        self.resume_from_checkpoint = resume_from_checkpoint
        self.metadata = metadata or {}
        self.torch_config = torch_config
        self.datasets = datasets or {}

    # ------------------------------------------------------------------------------------------
    def fit(self) -> Result:
        if "RANK" in os.environ and "WORLD_SIZE" in os.environ and not os.environ.get("RTDC_WORKER"):
            return self._fit_spmd()
        return self._fit_driver()

    def _params(self) -> Dict[str, Any]:
        return {"train_loop_config": {k: repr(v) for k, v in (self.train_loop_config or {}).items()},
                "scaling_config": asdict_shallow(self.scaling_config),
                "checkpoint_config": asdict_shallow(self.run_config.checkpoint_config)}

    # ------------------------------------------------------------------------------------------ driver mode
    def _fit_driver(self) -> Result:
        import cloudpickle
        from ..launch.group import WorkerGroup, worker_argv

        rc, sc = self.run_config, self.scaling_config
        storage = StorageContext(rc.storage_path, rc.name, storage_filesystem=rc.storage_filesystem)
        storage.make_dirs()
        book = _Bookkeeper(storage, rc, self._params())
        if sc.use_gpu:
            import torch
            if torch.cuda.is_available() and torch.cuda.device_count() < sc.num_workers * max(1, sc.num_gpus_per_worker):
                raise ValueError(f"ScalingConfig asks for {sc.num_workers} GPU workers but only {torch.cuda.device_count()} GPUs are visible")
        _log(rc.verbose, f"TorchTrainer: {sc.num_workers} worker(s), backend={sc.backend}, run dir {storage.trial_path}")

        max_failures = rc.failure_config.max_failures
        failures, attempt, next_index = 0, 0, 0
        start_ckpt = self.resume_from_checkpoint
        last_error: Optional[Exception] = None
        while True:
            ok, next_index, err = self._run_attempt(cloudpickle, WorkerGroup, worker_argv, storage, book, start_ckpt, next_index, attempt)
            if ok:
                last_error = None
                break
            last_error = err
            failures += 1
            if max_failures != -1 and failures > max_failures:
                break
            attempt += 1
            start_ckpt = book.manager.latest or start_ckpt
            _log(rc.verbose, f"worker group failed ({err}); restart {failures}/{max_failures if max_failures != -1 else 'inf'} "
                             f"from {start_ckpt.path if start_ckpt else 'scratch'}")
        result = book.result(last_error)
        if last_error is not None:
            raise TrainingFailedError(f"training failed after {failures} failure(s): {last_error}") from last_error
        _log(rc.verbose, f"training finished: {len(book.history)} report(s), latest checkpoint "
                         f"{result.checkpoint.path if result.checkpoint else None}")
        return result

    def _run_attempt(self, cloudpickle, WorkerGroup, worker_argv, storage, book, start_ckpt, next_index, attempt):
        rc, sc = self.run_config, self.scaling_config
        tmp = tempfile.mkdtemp(prefix="rtdc_run_")
        sock = os.path.join(tmp, "driver.sock")
        authkey = os.urandom(16)
        payload = os.path.join(tmp, "payload.pkl")
        with open(payload, "wb") as f:
            cloudpickle.dump({
                "fn": self.train_loop_per_worker, "config": self.train_loop_config, "metadata": self.metadata,
                "storage": storage, "use_gpu": sc.use_gpu, "backend": sc.backend, "start_checkpoint": start_ckpt,
                "start_index": next_index, "socket": sock, "authkey": authkey, "sys_path": list(sys.path),
                "experiment_name": storage.experiment_name, "trial_name": storage.trial_name,
                "heartbeat_s": rc.worker_heartbeat_s if rc.worker_report_timeout_s else 0.0,
            }, f)
        server = DriverServer(sock, sc.num_workers, authkey)
        # workers must be able to import whatever module the training function lives in
        py_path = os.pathsep.join([p for p in sys.path if p] + [os.environ.get("PYTHONPATH", "")]).strip(os.pathsep)
        group = WorkerGroup(sc.num_workers, worker_argv(payload), env={"RTDC_ATTEMPT": str(attempt), "PYTHONPATH": py_path})
        group.start()
        pending: Dict[int, List[dict]] = {}
        done_ranks, hello_ranks = set(), set()
        error: Optional[Exception] = None
        t_start = time.time()
        last_progress = time.time()
        try:
            while True:
                try:
                    ev = server.events.get(timeout=0.05)
                except Exception:
                    ev = None
                if ev is not None:
                    last_progress = time.time()
                    kind = ev["type"]
                    if kind == "hello":
                        hello_ranks.add(ev["rank"])
                    elif kind == "heartbeat":
                        pass                                   # liveness only: last_progress was refreshed above
                    elif kind == "report":
                        idx = ev["index"]
                        pending.setdefault(idx, []).append(ev)
                        if len(pending[idx]) == sc.num_workers:
                            has_ckpt = book.complete(idx, pending.pop(idx))
                            next_index = idx + 1
                            server.broadcast({"type": "ack", "index": idx, "has_checkpoint": has_ckpt})
                    elif kind == "done":
                        done_ranks.add(ev["rank"])
                    elif kind == "error":
                        error = RuntimeError(f"worker rank {ev['rank']} raised:\n{ev.get('traceback', '')}")
                        break
                    elif kind == "eof" and ev["rank"] not in done_ranks:
                        error = RuntimeError(f"worker rank {ev['rank']} disconnected unexpectedly")
                        break
                code = group.poll()
                if code is not None and code != 0:
                    # drain a possible error message that explains the exit
                    t_end = time.time() + 0.5
                    while time.time() < t_end and error is None:
                        try:
                            ev = server.events.get(timeout=0.05)
                        except Exception:
                            continue
                        if ev["type"] == "error":
                            error = RuntimeError(f"worker rank {ev['rank']} raised:\n{ev.get('traceback', '')}")
                    error = error or RuntimeError(f"a worker process exited with status {code}")
                    break
                if len(done_ranks) == sc.num_workers and code == 0:
                    break
                if len(hello_ranks) < sc.num_workers and time.time() - t_start > rc.worker_start_timeout_s:
                    error = TimeoutError(f"only {len(hello_ranks)}/{sc.num_workers} workers started within {rc.worker_start_timeout_s}s")
                    break
                if rc.worker_report_timeout_s and time.time() - last_progress > rc.worker_report_timeout_s:
                    error = TimeoutError(f"no worker activity for {rc.worker_report_timeout_s}s")
                    break
        finally:
            if error is not None:
                server.broadcast({"type": "abort"})
            group.shutdown(grace_s=0.5 if error is not None else 10.0)
            server.close()
            shutil.rmtree(tmp, ignore_errors=True)
        return error is None, next_index, error

    # ------------------------------------------------------------------------------------------ SPMD mode
    def _fit_spmd(self) -> Result:
        import torch
        import torch.distributed as dist

        rc, sc = self.run_config, self.scaling_config
        rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
        local_rank = int(os.environ.get("LOCAL_RANK", rank))
        if sc.num_workers != world:
            raise ValueError(f"ScalingConfig(num_workers={sc.num_workers}) but the launch has WORLD_SIZE={world}")
        if sc.use_gpu:
            from ..utils.affinity import bind_rank_cores
            bind_rank_cores(local_rank, int(os.environ.get("LOCAL_WORLD_SIZE", world)))   # the native launcher pins its workers too
        created_pg = False
        if world > 1 and not dist.is_initialized():
            kwargs = {}
            if sc.use_gpu:
                torch.cuda.set_device(local_rank)
                kwargs["device_id"] = torch.device("cuda", local_rank)
            dist.init_process_group(backend=sc.backend, **kwargs)
            created_pg = True
        elif sc.use_gpu:
            torch.cuda.set_device(local_rank)
        names = [rc.name, None]
        if rank == 0:
            s0 = StorageContext(rc.storage_path, rc.name, storage_filesystem=rc.storage_filesystem)
            names = [s0.experiment_name, s0.trial_name]
        if world > 1:
            dist.broadcast_object_list(names, src=0)
        storage = StorageContext(rc.storage_path, names[0], names[1], storage_filesystem=rc.storage_filesystem)
        book = None
        if rank == 0:
            storage.make_dirs()
            book = _Bookkeeper(storage, rc, self._params())
        if world > 1:
            dist.barrier()
        link = SpmdLink(rank, world, book.complete if book else None)
        ctx = TrainContext(rank, world, local_rank, int(os.environ.get("LOCAL_WORLD_SIZE", world)),
                           int(os.environ.get("NODE_RANK", 0)), storage.experiment_name, storage.trial_name,
                           storage.trial_path, storage, sc.use_gpu, self.metadata)
        session = _Session(ctx, link, self.resume_from_checkpoint)
        _set_session(session)
        error = None
        try:
            _call_train_fn(self.train_loop_per_worker, self.train_loop_config)
            session.drain()
            from ..data.loader import dump_timing
            dump_timing(rank)
        except Exception as e:     # surface through Result like the driver mode does
            error = e
            raise
        finally:
            _set_session(None)
            if error is None:
                link.close()
            payload = [book.result() if book else None]
            if error is None and world > 1 and dist.is_initialized():
                dist.broadcast_object_list(payload, src=0)      # every rank returns the Result, before the group goes away
            if created_pg and error is None:
                dist.barrier()
                dist.destroy_process_group()
        return payload[0] if payload[0] is not None else Result(metrics=None, checkpoint=None, path=storage.trial_path)
