"""Asynchronous checkpoint snapshots (K11 in SURVEY.md §2.5).

The reference serialises its checkpoint synchronously, twice, on every rank
(``torch.save`` at ``/root/reference/my_ray_module.py:179-201``): every tensor is copied device->host
inside the pickler while the GPU idles. Here a checkpoint is taken in three decoupled stages:

1. **device snapshot** on the training stream: every CUDA tensor of the state is gathered into one
   flat device staging buffer (HBM is 180 GB; a second copy of the state is cheap and the copy runs
   at HBM speed), so the optimizer may overwrite the live tensors immediately afterwards;
2. **D2H** of that staging buffer into pinned host memory on a side stream, overlapping the next
   training steps;
3. **serialise + publish** on a host thread: ``torch.save`` of CPU views (the stock zip/pickle
   format, loadable with ``torch.load(weights_only=True)`` exactly like the reference's files),
   written to a temp name and renamed into place.

``save()`` returns a ``PendingCheckpoint`` that ``train.report(..., checkpoint=pending)`` accepts; the
report then completes in the background as well.
"""
from __future__ import annotations

import concurrent.futures as _fut
import os
import tempfile
import threading
from typing import Any, Dict, List, Optional, Tuple

import torch

_ALIGN = 256
_registry: List["AsyncCheckpointer"] = []
_registry_lock = threading.Lock()


def _align(n: int) -> int:
    return (n + _ALIGN - 1) // _ALIGN * _ALIGN


FAST_WRITER_MIN_BYTES = 64 << 20     # states at least this big go through the parallel writer (ckpt/fast_zip.py)


def atomic_torch_save(obj: Any, path: str, parallel: Optional[bool] = None) -> None:
    """``torch.save`` to a temp file in the same directory, then ``rename`` (readers never see a
    partial file; concurrent writers of the same name cannot interleave). States of 64 MB and more are written by the
    parallel archive writer (same entries and contents, ``torch.load``-able; ``parallel=False`` or
    ``RTDC_CKPT_WRITER=torch`` keeps stock ``torch.save``), which falls back to ``torch.save`` if it cannot handle the object."""
    d = os.path.dirname(os.path.abspath(path)) or "."
    os.makedirs(d, exist_ok=True)
    fd, tmp = tempfile.mkstemp(prefix="." + os.path.basename(path) + ".", suffix=".tmp", dir=d)
    os.close(fd)
    try:
        done = False
        if parallel is None:
            from .fast_zip import state_bytes
            parallel = os.environ.get("RTDC_CKPT_WRITER", "") != "torch" and state_bytes(obj) >= FAST_WRITER_MIN_BYTES
        if parallel:
            try:
                from . import fast_zip
                fast_zip.save(obj, tmp, archive_name=os.path.splitext(os.path.basename(path))[0])
                done = True
            except Exception as e:                       # private PyTorch hook moved, exotic object, ...: the stock path always works
                import warnings
                warnings.warn(f"parallel checkpoint writer unavailable ({type(e).__name__}: {e}); using torch.save")
        if not done:
            # the archive's root folder is the file stem, as for a direct torch.save(obj, path)
            _save_with_stem(obj, tmp, os.path.basename(path))
        os.replace(tmp, path)
    except BaseException:
        if os.path.exists(tmp):
            os.unlink(tmp)
        raise


def _save_with_stem(obj: Any, tmp_path: str, final_name: str) -> None:
    """torch.save names the zip's root directory after the *file name it was given*. Saving under
    the final name inside a private temp directory keeps the archive identical to what a direct
    ``torch.save(obj, final_name)`` produces (entries ``<stem>/data.pkl`` ...)."""
    d = tempfile.mkdtemp(prefix=".rtdc_save_", dir=os.path.dirname(tmp_path))
    try:
        p = os.path.join(d, final_name)
        torch.save(obj, p)
        os.replace(p, tmp_path)
    finally:
        try:
            os.rmdir(d)
        except OSError:
            pass


def _link_or_copy(src: str, dst: str) -> None:
    tmp = dst + f".{os.getpid()}.tmp"
    try:
        os.link(src, tmp)
    except OSError:
        import shutil
        shutil.copyfile(src, tmp)
    os.replace(tmp, dst)


class PendingCheckpoint:
    """Handle to a checkpoint directory whose files are still being written."""

    def __init__(self, directory: str, futures: List[_fut.Future]):
        self.directory = directory
        self._futures = futures

    def result(self, timeout: Optional[float] = None):
        from ..train.checkpoint import Checkpoint
        for f in self._futures:
            f.result(timeout)
        return Checkpoint.from_directory(self.directory)

    def done(self) -> bool:
        return all(f.done() for f in self._futures)

    @property
    def path(self) -> str:
        return self.directory


class _Slot:
    def __init__(self):
        self.dev: Optional[torch.Tensor] = None
        self.host: Optional[torch.Tensor] = None
        self.busy: Optional[_fut.Future] = None


class AsyncCheckpointer:
    """Double-buffered snapshot -> pinned D2H -> writer-thread pipeline for one device."""

    def __init__(self, device: Optional[torch.device] = None, slots: int = 2):
        self.device = torch.device(device) if device is not None else (
            torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else torch.device("cpu"))
        self.is_cuda = self.device.type == "cuda"
        self.stream = torch.cuda.Stream(self.device) if self.is_cuda else None
        self.slots = [_Slot() for _ in range(max(1, slots))]
        self._next = 0
        self.pool = _fut.ThreadPoolExecutor(max_workers=1, thread_name_prefix="rtdc-ckpt")
        self._inflight: List[_fut.Future] = []
        self.stats = {"snapshots": 0, "bytes": 0, "critical_path_s": 0.0}
        with _registry_lock:
            _registry.append(self)

    # -------------------------------------------------------------------------------------
    @staticmethod
    def _collect(obj, out: List[torch.Tensor]):
        if isinstance(obj, torch.Tensor):
            out.append(obj)
        elif isinstance(obj, dict):
            for v in obj.values():
                AsyncCheckpointer._collect(v, out)
        elif isinstance(obj, (list, tuple)):
            for v in obj:
                AsyncCheckpointer._collect(v, out)

    @staticmethod
    def _rebuild(obj, mapping: Dict[int, torch.Tensor]):
        if isinstance(obj, torch.Tensor):
            return mapping[id(obj)]
        if isinstance(obj, dict):
            return type(obj)((k, AsyncCheckpointer._rebuild(v, mapping)) for k, v in obj.items())
        if isinstance(obj, list):
            return [AsyncCheckpointer._rebuild(v, mapping) for v in obj]
        if isinstance(obj, tuple):
            return tuple(AsyncCheckpointer._rebuild(v, mapping) for v in obj)
        return obj

    def _acquire_slot(self, nbytes: int) -> _Slot:
        slot = self.slots[self._next]
        self._next = (self._next + 1) % len(self.slots)
        if slot.busy is not None:
            slot.busy.result()          # both buffers in flight: wait for the older write
            slot.busy = None
        if slot.host is None or slot.host.numel() < nbytes:
            cap = max(nbytes, 1 << 20)
            slot.host = torch.empty(cap, dtype=torch.uint8, pin_memory=self.is_cuda)
            if self.is_cuda:
                slot.dev = torch.empty(cap, dtype=torch.uint8, device=self.device)
        return slot

    def snapshot(self, state: Any):
        """Stage ``state`` (nested dict/list of tensors and plain values) and return
        ``(cpu_state_factory, slot)`` where the factory blocks until the bytes are on the host."""
        import time
        t0 = time.perf_counter()
        tensors: List[torch.Tensor] = []
        self._collect(state, tensors)
        uniq: Dict[int, torch.Tensor] = {}
        for t in tensors:
            uniq.setdefault(id(t), t)
        # Tensors that are views of one flat arena (the engine's parameter / momentum arenas, a DDP wrapper's) are snapshotted
        # with ONE copy of the arena range they cover instead of one copy per tensor: a checkpoint of the reference MLP is two
        # device copies, not twelve.
        offsets: Dict[int, Tuple[int, int]] = {}
        total = 0
        groups: Dict[int, List[int]] = {}
        for k, t in uniq.items():
            if t.is_cuda and t.is_contiguous() and t.numel():
                groups.setdefault(t.untyped_storage().data_ptr(), []).append(k)
        ranges = []                                   # (staging offset, source byte tensor view) per grouped arena range
        grouped = set()
        for keys in groups.values():
            if len(keys) < 2:
                continue
            ts = [uniq[k] for k in keys]
            lo = min(t.data_ptr() for t in ts)
            hi = max(t.data_ptr() + t.numel() * t.element_size() for t in ts)
            covered = sum(t.numel() * t.element_size() for t in ts)
            if covered * 2 < hi - lo:                 # sparse use of a big storage: copy the tensors one by one
                continue
            base = ts[0].untyped_storage()
            start = lo - base.data_ptr()
            src = torch.empty(0, dtype=torch.uint8, device=ts[0].device).set_(base, start, (hi - lo,))
            ranges.append((total, src))
            for k, t in zip(keys, ts):
                offsets[k] = (total + (t.data_ptr() - lo), t.numel() * t.element_size())
                grouped.add(k)
            total += _align(hi - lo)
        for k, t in uniq.items():
            if k in grouped:
                continue
            n = t.numel() * t.element_size()
            offsets[k] = (total, n)
            total += _align(n)
        t_plan = time.perf_counter()
        slot = self._acquire_slot(total)
        t_slot = time.perf_counter()
        done_event = None
        if self.is_cuda:
            main = torch.cuda.current_stream(self.device)
            for off, src in ranges:
                slot.dev[off:off + src.numel()].copy_(src, non_blocking=True)          # D2D at HBM speed, one per arena
            for k, t in uniq.items():
                off, n = offsets[k]
                if n == 0 or k in grouped:
                    continue
                dst = slot.dev[off:off + n].view(t.dtype)
                if t.is_cuda:
                    dst.copy_(t.detach().reshape(-1), non_blocking=True)           # D2D at HBM speed
                else:
                    slot.host[off:off + n].view(t.dtype).copy_(t.detach().reshape(-1))
            ev = torch.cuda.Event()
            ev.record(main)
            with torch.cuda.stream(self.stream):
                self.stream.wait_event(ev)
                has_cuda = any(t.is_cuda for t in uniq.values())
                if has_cuda and total:
                    if all(t.is_cuda for t in uniq.values()):
                        slot.host[:total].copy_(slot.dev[:total], non_blocking=True)   # one pinned D2H
                    else:
                        for k, t in uniq.items():
                            off, n = offsets[k]
                            if t.is_cuda and n:
                                slot.host[off:off + n].copy_(slot.dev[off:off + n], non_blocking=True)
                done_event = torch.cuda.Event()
                done_event.record(self.stream)
        else:
            for k, t in uniq.items():
                off, n = offsets[k]
                if n:
                    slot.host[off:off + n].view(t.dtype).copy_(t.detach().reshape(-1))

        def cpu_state(alias: bool = False):
            """The state as CPU tensors, each owning a storage of exactly its own size (as in a stock state_dict). By default
            they are clones; ``alias=True`` gives zero-copy views of the pinned snapshot buffer, valid until the slot is
            reused (what ``save`` hands to the writer: no staging copy of a gigabyte-sized state)."""
            if done_event is not None:
                done_event.synchronize()
            mapping = {}
            host_np = slot.host.numpy() if alias else None
            for k, t in uniq.items():
                off, n = offsets[k]
                if alias and n:
                    mapping[k] = torch.frombuffer(host_np, dtype=t.dtype, count=t.numel(), offset=off).reshape(t.shape)
                else:
                    mapping[k] = slot.host[off:off + n].view(t.dtype).reshape(t.shape).clone()
            return self._rebuild(state, mapping)

        self.stats["snapshots"] += 1
        self.stats["bytes"] += total
        self.stats["critical_path_s"] += time.perf_counter() - t0
        if os.environ.get("RTDC_PROFILE_EPOCHS"):
            print(f"[rtdc] snapshot: plan {(t_plan - t0) * 1e3:.2f} ms, wait for a free slot {(t_slot - t_plan) * 1e3:.2f} ms, enqueue copies "
                  f"{(time.perf_counter() - t_slot) * 1e3:.2f} ms", flush=True)
        return cpu_state, slot

    def save(self, state: Any, directory: str, filenames) -> PendingCheckpoint:
        """Snapshot ``state`` now; write it to ``directory/<name>`` for each name in the background."""
        if isinstance(filenames, str):
            filenames = [filenames]
        os.makedirs(directory, exist_ok=True)
        cpu_state, slot = self.snapshot(state)
        from .fast_zip import state_bytes
        big = state_bytes(state) >= FAST_WRITER_MIN_BYTES

        def job():
            import time
            t0 = time.time()
            # big states are handed to the writer as views of the pinned snapshot buffer (no staging copy of a gigabyte);
            # obj is dropped before this job ends and the slot stays busy until then. Small ones keep the cloning path.
            obj = cpu_state(alias=big)
            t1 = time.time()
            first = os.path.join(directory, filenames[0])
            atomic_torch_save(obj, first)
            if os.environ.get("RTDC_PROFILE_EPOCHS"):
                print(f"[rtdc] checkpoint writer: D2H wait + host copies {(t1 - t0) * 1e3:.2f} ms, torch.save {(time.time() - t1) * 1e3:.2f} ms "
                      f"({len(filenames)} file(s))", flush=True)
            for name in filenames[1:]:
                # further names of the same state (best_model.pt next to latest_model.pt): serialise once more only when the
                # archive's root folder has to differ (torch names it after the file stem), else link the bytes
                if os.path.splitext(name)[0] == os.path.splitext(filenames[0])[0]:
                    _link_or_copy(first, os.path.join(directory, name))
                else:
                    atomic_torch_save(obj, os.path.join(directory, name))
            return directory

        fut = self.pool.submit(job)
        slot.busy = fut
        self._inflight = [f for f in self._inflight if not f.done()] + [fut]
        return PendingCheckpoint(directory, [fut])

    def wait(self):
        for f in list(self._inflight):
            f.result()
        self._inflight = []

    def close(self):
        self.wait()
        self.pool.shutdown(wait=True)


def wait_all():
    """Block until every AsyncCheckpointer of this process has published its files."""
    with _registry_lock:
        items = list(_registry)
    for c in items:
        c.wait()
