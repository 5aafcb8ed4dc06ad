"""Sharded sampling and host->device batch movement.

* ``DistributedSampler`` / ``shard_data_loader``: what ``ray.train.torch.prepare_data_loader`` does to
  the reference's loaders (``/root/reference/my_ray_module.py:128-129``) — a per-rank, per-epoch
  (``set_epoch``) shard of the dataset, same index arithmetic as ``torch.utils.data.DistributedSampler``
  (pad to a multiple of the world size by wrapping around, then stride by rank).
* ``DeviceLoader``: moves every batch to the worker's device through pinned memory on a copy stream,
  one batch ahead (K9).
* ``U8BatchLoader``: the B200-first input path for image data that is natively uint8 (FashionMNIST):
  the dataset stays in pinned host memory as bytes, a batch is gathered by native threads, crosses
  PCIe as uint8 (4x fewer bytes than the reference's float32 batches) and is normalised on the GPU by
  ``ops.u8_normalize``.
"""
from __future__ import annotations

import math
import threading
from typing import Iterator, Optional

import numpy as np
import torch
from torch.utils.data import DataLoader, RandomSampler, Sampler


class DistributedSampler(Sampler):
    def __init__(self, dataset_len: int, num_replicas: int, rank: int, shuffle: bool = True, seed: int = 0,
                 drop_last: bool = False):
        if rank < 0 or rank >= num_replicas:
            raise ValueError(f"rank {rank} out of range for {num_replicas} replicas")
        self.n, self.num_replicas, self.rank = int(dataset_len), num_replicas, rank
        self.shuffle, self.seed, self.drop_last, self.epoch = shuffle, seed, drop_last, 0
        if drop_last and self.n % num_replicas:
            self.num_samples = math.ceil((self.n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(self.n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    def indices(self) -> list:
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g).tolist()
        else:
            idx = list(range(self.n))
        if not self.drop_last:
            pad = self.total_size - len(idx)
            if pad > 0:
                idx += (idx * math.ceil(pad / max(1, len(idx))))[:pad]
        else:
            idx = idx[:self.total_size]
        return idx[self.rank:self.total_size:self.num_replicas]

    def __iter__(self) -> Iterator[int]:
        return iter(self.indices())

    def __len__(self) -> int:
        return self.num_samples


def shard_data_loader(loader: DataLoader, world: int, rank: int) -> DataLoader:
    """Rebuild ``loader`` around a DistributedSampler (shuffle iff the original sampler was random)."""
    shuffle = isinstance(getattr(loader, "sampler", None), RandomSampler)
    sampler = DistributedSampler(len(loader.dataset), world, rank, shuffle=shuffle)
    return DataLoader(loader.dataset, batch_size=loader.batch_size, sampler=sampler, num_workers=loader.num_workers,
                      collate_fn=loader.collate_fn, pin_memory=loader.pin_memory, drop_last=loader.drop_last,
                      timeout=loader.timeout, worker_init_fn=loader.worker_init_fn,
                      persistent_workers=getattr(loader, "persistent_workers", False) and loader.num_workers > 0)


def _to_device(obj, device, non_blocking):
    if isinstance(obj, torch.Tensor):
        if device.type == "cuda" and not obj.is_cuda and not obj.is_pinned() and non_blocking:
            obj = obj.pin_memory()
        return obj.to(device, non_blocking=non_blocking)
    if isinstance(obj, (list, tuple)):
        return type(obj)(_to_device(o, device, non_blocking) for o in obj)
    if isinstance(obj, dict):
        return {k: _to_device(v, device, non_blocking) for k, v in obj.items()}
    return obj


class DeviceLoader:
    """Iterates ``loader`` with every batch already on ``device``; on CUDA the next batch's H2D copy
    runs on a side stream while the current one is consumed."""

    def __init__(self, loader, device: torch.device, prefetch: bool = True, timed: bool = False):
        self.loader = loader
        self.timed = timed
        self.device = torch.device(device)
        self.prefetch = prefetch and self.device.type == "cuda"
        self.stream = torch.cuda.Stream(self.device) if self.prefetch else None

    @property
    def sampler(self):
        return self.loader.sampler

    @property
    def dataset(self):
        return self.loader.dataset

    @property
    def batch_size(self):
        return self.loader.batch_size

    def __len__(self):
        return len(self.loader)

    def __iter__(self):
        if not self.prefetch:
            for batch in self.loader:
                moved = _to_device(batch, self.device, False)
                if self.timed:
                    _stamp("step_begin", self.device)
                yield moved
                if self.timed:
                    _stamp("step_end", self.device)
            return
        it = iter(self.loader)

        def stage():
            try:
                b = next(it)
            except StopIteration:
                return None, None
            with torch.cuda.stream(self.stream):
                moved = _to_device(b, self.device, True)
                ev = torch.cuda.Event()
                ev.record(self.stream)
            return moved, ev

        nxt, ev = stage()
        while nxt is not None:
            torch.cuda.current_stream(self.device).wait_event(ev)
            cur = nxt
            _record_stream(cur, torch.cuda.current_stream(self.device))
            nxt, ev = stage()
            if self.timed:
                _stamp("step_begin", self.device)
            yield cur
            if self.timed:
                _stamp("step_end", self.device)


def _record_stream(obj, stream):
    if isinstance(obj, torch.Tensor):
        if obj.is_cuda:
            obj.record_stream(stream)
    elif isinstance(obj, (list, tuple)):
        for o in obj:
            _record_stream(o, stream)
    elif isinstance(obj, dict):
        for o in obj.values():
            _record_stream(o, stream)


class U8Images(torch.Tensor):
    """A batch of uint8 images that still knows its normalisation.

    The bytes are what crossed PCIe (``raw``: ``uint8 [B, ...]``, 4x fewer than float32). Code that understands it — the
    engine-backed model of ``train.torch.prepare_model`` — consumes the bytes directly and folds ``(x/255 - mean)/std`` into
    its first GEMM. Any other torch operation (``nn.Flatten``, ``.float()``, ``.to()``, arithmetic ...) first turns the
    batch into the normalised ``float32`` tensor the reference's ``ToTensor() + Normalize`` pipeline would have produced
    (``/root/reference/my_ray_module.py:38``; one ``ops.u8_normalize`` launch, cached), so generic models work unchanged.
    """

    @staticmethod
    def wrap(raw: torch.Tensor, mean: float, std: float) -> "U8Images":
        t = torch.Tensor._make_subclass(U8Images, raw)
        t.norm = (float(mean), float(std))
        t._f32 = None
        return t

    @property
    def raw(self) -> torch.Tensor:
        with torch._C.DisableTorchFunctionSubclass():
            return self.as_subclass(torch.Tensor)

    def normalized(self) -> torch.Tensor:
        if self._f32 is None:
            r = self.raw
            if r.is_cuda:
                from ..ops import u8_normalize
                self._f32 = u8_normalize(r.contiguous(), mean=self.norm[0], std=self.norm[1])
            else:
                self._f32 = (r.to(torch.float32) / 255.0 - self.norm[0]) / self.norm[1]
        return self._f32

    # attribute / bookkeeping calls answered by the byte tensor itself; everything else sees normalised float32
    _PASS = frozenset({"__get__", "size", "dim", "numel", "stride", "is_contiguous", "record_stream", "__len__", "data_ptr",
                       "element_size", "is_pinned", "storage_offset", "untyped_storage", "__repr__", "__str__", "__format__",
                       "__hash__", "_is_view", "get_device"})

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        kwargs = kwargs or {}
        if getattr(func, "__name__", None) in cls._PASS:
            with torch._C.DisableTorchFunctionSubclass():
                return func(*args, **kwargs)
        conv = lambda v: v.normalized() if isinstance(v, U8Images) else v
        with torch._C.DisableTorchFunctionSubclass():
            return func(*[conv(a) for a in args], **{k: conv(v) for k, v in kwargs.items()})


# Step stopwatch of the training loader, the same one the benchmark's reference arm carries in its ``ray.train.torch`` stand-in
# (baseline/ray_standin/ray/train/torch.py): a CUDA event + host time when a batch is handed to the loop ("step_begin") and
# when the loop comes back for the next one ("step_end"). Off unless RTDC_STEP_TIMING is set.
TIMING = {"records": [], "loaders": 0, "enabled": None, "extra": {}}


def _timing_enabled() -> bool:
    if TIMING["enabled"] is None:
        import os
        TIMING["enabled"] = bool(os.environ.get("RTDC_STEP_TIMING"))
    return TIMING["enabled"]


def _claim_timed_loader() -> bool:
    """The first loader prepared in a worker (the reference prepares the training loader first, my_ray_module.py:128)."""
    if not _timing_enabled():
        return False
    first = TIMING["loaders"] == 0
    TIMING["loaders"] += 1
    return first


def _stamp(kind: str, device: torch.device):
    import time
    ev = None
    if device.type == "cuda":
        ev = torch.cuda.Event(enable_timing=True)
        ev.record(torch.cuda.current_stream(device))
    TIMING["records"].append((kind, time.perf_counter(), ev))


def timing_summary(skip: int = 0):
    """Per-step device intervals (ms) and host stamps (s) of the timed loader, after a device synchronize."""
    rec = TIMING["records"]
    begins = [r for r in rec if r[0] == "step_begin"]
    ends = [r for r in rec if r[0] == "step_end"]
    n = min(len(begins), len(ends))
    if n and begins[0][2] is not None:
        torch.cuda.synchronize()
        dev = [begins[i][2].elapsed_time(ends[i][2]) for i in range(n)]
        since_first = [begins[0][2].elapsed_time(ends[i][2]) for i in range(n)]
        begin_since_first = [begins[0][2].elapsed_time(begins[i][2]) for i in range(n)]
    else:
        dev = [(ends[i][1] - begins[i][1]) * 1e3 for i in range(n)]
        since_first = [(ends[i][1] - begins[0][1]) * 1e3 for i in range(n)]
        begin_since_first = [(begins[i][1] - begins[0][1]) * 1e3 for i in range(n)]
    return {"extra": dict(TIMING["extra"]), "steps": n, "device_ms": dev, "end_ms_since_first_begin": since_first, "begin_ms_since_first_begin": begin_since_first,
            "host_begin_s": [b[1] for b in begins[:n]], "host_end_s": [e[1] for e in ends[:n]]}


def dump_timing(rank: int):
    """Write this worker's step stopwatch to ``$RTDC_STEP_TIMING.<rank>`` (JSON) when the variable names a path."""
    import json
    import os
    path = os.environ.get("RTDC_STEP_TIMING")
    if not path or path in ("1", "true") or not TIMING["records"]:
        return
    tmp = f"{path}.{rank}.tmp"
    with open(tmp, "w") as f:
        json.dump(timing_summary(), f)
    os.replace(tmp, f"{path}.{rank}")


class U8BatchLoader:
    """Epoch iterator over a uint8 image set held in host memory, built for PCIe + HBM rather than for
    a Python collate loop.

    * The shuffle is done once per epoch, off the critical path: a background thread gathers this
      rank's shard of the epoch's permutation into a pinned buffer (``index_select`` on all host cores)
      while the previous epoch is still training (two pinned buffers, ping-pong).
    * Every step is then ONE contiguous ``cudaMemcpyAsync`` of uint8 pixels plus one of uint8 labels on a
      copy stream — 4x fewer pixel bytes than the reference's float32 batches and 8x fewer label bytes than int64
      (my_ray_module.py:38, 128) — issued ``lookahead`` steps ahead of the consumer, into a ring of device buffers;
      labels are widened to int64 on the device.
    * Normalisation (ToTensor + Normalize) happens on the GPU: the consumer takes the uint8 batch (the fused
      engine normalises inside its graph); with ``wrap=True`` batches are :class:`U8Images` (image-shaped, normalise
      themselves on first generic use); with ``normalize=True`` this loader runs ``ops.u8_normalize`` and yields float32.

    Yields ``(x, labels)`` on the device; ``x`` is ``uint8 [B, features]``, ``U8Images [B, *image_shape]`` or
    ``float32 [B, features]``; ``labels`` is ``int64 [B]``.
    """

    def __init__(self, images_u8: torch.Tensor, labels: torch.Tensor, batch_size: int, device, *, shuffle: bool = True,
                 world: int = 1, rank: int = 0, seed: int = 0, drop_last: bool = True, normalize: bool = False,
                 mean: float = 0.5, std: float = 0.5, lookahead: int = 2, prefetch_next_epoch: bool = True,
                 wrap: bool = False, timed: bool = False):
        assert images_u8.dtype == torch.uint8 and labels.dtype in (torch.int64, torch.uint8)
        self.prefetch_next_epoch = prefetch_next_epoch
        self.image_shape = tuple(images_u8.shape[1:]) if images_u8.dim() > 2 else (images_u8.shape[1],)
        if len(self.image_shape) == 2:
            self.image_shape = (1,) + self.image_shape          # ToTensor's channel dimension
        self.images = images_u8.reshape(images_u8.shape[0], -1).contiguous()
        # labels cross PCIe as bytes when they fit (10 classes do)
        self.label_dtype = torch.uint8 if (labels.dtype == torch.uint8 or (labels.numel() and 0 <= int(labels.min()) and int(labels.max()) < 256)) else torch.int64
        self.labels = labels.to(self.label_dtype).contiguous()
        self.row = self.images.shape[1]
        self.batch_size = int(batch_size)
        self.device = torch.device(device)
        self.sampler = DistributedSampler(self.images.shape[0], world, rank, shuffle=shuffle, seed=seed)
        self.drop_last = drop_last
        self.normalize, self.mean, self.std = normalize, mean, std
        self.wrap, self.timed = wrap, timed
        self.cuda = self.device.type == "cuda"
        self.lookahead = max(1, int(lookahead))
        n = len(self.sampler)
        self._host_x = [torch.empty(n, self.row, dtype=torch.uint8, pin_memory=self.cuda) for _ in range(2)]
        self._host_y = [torch.empty(n, dtype=self.label_dtype, pin_memory=self.cuda) for _ in range(2)]
        self._prepared = {}          # epoch -> (slot, thread)
        self._epoch_iter = 0
        if self.cuda:
            ring = self.lookahead + 2
            self._dev_x = [torch.empty(self.batch_size, self.row, dtype=torch.uint8, device=self.device) for _ in range(ring)]
            self._dev_y = [torch.empty(self.batch_size, dtype=torch.int64, device=self.device) for _ in range(ring)]
            self._dev_y8 = ([torch.empty(self.batch_size, dtype=torch.uint8, device=self.device) for _ in range(ring)]
                            if self.label_dtype == torch.uint8 else None)
            self._dev_f = ([torch.empty(self.batch_size, self.row, dtype=torch.float32, device=self.device) for _ in range(ring)]
                           if normalize else None)
            self._free_ev = [None] * ring
            self.copy_stream = torch.cuda.Stream(self.device)
        self.h2d_bytes_per_batch = self.batch_size * (self.row + (1 if self.label_dtype == torch.uint8 else 8))
        if timed:
            TIMING["extra"]["h2d_bytes_per_step"] = self.h2d_bytes_per_batch
        self._prepare(self.sampler.epoch)

    @property
    def dataset(self):
        return self.images

    def __len__(self):
        n = len(self.sampler)
        return n // self.batch_size if self.drop_last else math.ceil(n / self.batch_size)

    # ---------------------------------------------------------------- epoch permutation (background)
    def _prepare(self, epoch: int):
        if epoch in self._prepared:
            return
        slot = epoch % 2
        old_epoch = self.sampler.epoch
        self.sampler.set_epoch(epoch)
        idx = torch.as_tensor(self.sampler.indices(), dtype=torch.int64)
        self.sampler.set_epoch(old_epoch)

        def work():
            torch.index_select(self.images, 0, idx, out=self._host_x[slot])
            torch.index_select(self.labels, 0, idx, out=self._host_y[slot])

        th = threading.Thread(target=work, name="rtdc-epoch-permute", daemon=True)
        th.start()
        self._prepared[epoch] = (slot, th)

    def _emit(self, xb: torch.Tensor, yb: torch.Tensor):
        if self.normalize:
            return xb, yb
        if self.wrap:
            return U8Images.wrap(xb.view(xb.shape[0], *self.image_shape), self.mean, self.std), yb
        return xb, yb

    def __iter__(self):
        epoch = self.sampler.epoch
        self._prepare(epoch)
        slot, th = self._prepared.pop(epoch)
        th.join()
        for e in list(self._prepared):          # stale preparations (epoch set backwards)
            if e != epoch + 1:
                self._prepared.pop(e)[1].join()
        hx, hy = self._host_x[slot], self._host_y[slot]
        nb, bs = len(self), self.batch_size
        total = len(self.sampler)
        if nb > 1 and self.prefetch_next_epoch:
            self._prepare(epoch + 1)            # next epoch's shuffle overlaps this epoch's training
        if not self.cuda:
            for i in range(nb):
                x = hx[i * bs:min(total, (i + 1) * bs)]
                y = hy[i * bs:min(total, (i + 1) * bs)].to(torch.int64)
                if self.normalize:
                    x = (x.to(torch.float32) / 255.0 - self.mean) / self.std
                else:
                    x = x.clone()
                if self.timed:
                    _stamp("step_begin", self.device)
                yield self._emit(x, y)
                if self.timed:
                    _stamp("step_end", self.device)
            return
        ring = len(self._dev_x)
        inflight = []                            # (buffer index, rows, copy-done event)
        issued = 0

        def issue():
            nonlocal issued
            i = issued
            bi = i % ring
            lo, hi = i * bs, min(total, (i + 1) * bs)
            if self._free_ev[bi] is not None:
                self.copy_stream.wait_event(self._free_ev[bi])      # consumer is done with this device buffer
            with torch.cuda.stream(self.copy_stream):
                self._dev_x[bi][:hi - lo].copy_(hx[lo:hi], non_blocking=True)
                if self._dev_y8 is not None:
                    self._dev_y8[bi][:hi - lo].copy_(hy[lo:hi], non_blocking=True)
                    self._dev_y[bi][:hi - lo].copy_(self._dev_y8[bi][:hi - lo])        # widen on the device
                else:
                    self._dev_y[bi][:hi - lo].copy_(hy[lo:hi], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.copy_stream)
            inflight.append((bi, hi - lo, ev))
            issued += 1

        while issued < min(nb, self.lookahead):
            issue()
        for _ in range(nb):
            bi, rows, ev = inflight.pop(0)
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)
            if issued < nb:
                issue()                          # keep the copy stream `lookahead` batches ahead
            xb = self._dev_x[bi][:rows]
            if self.normalize:
                from ..ops import u8_normalize
                xb = u8_normalize(xb, mean=self.mean, std=self.std, out=self._dev_f[bi][:rows])
            if self.timed:
                _stamp("step_begin", self.device)
            yield self._emit(xb, self._dev_y[bi][:rows])
            if self.timed:
                _stamp("step_end", self.device)
            done = torch.cuda.Event()
            done.record(torch.cuda.current_stream(self.device))
            self._free_ev[bi] = done
        # the pinned epoch buffer may be rewritten (epoch + 2) only after its last H2D has finished
        self.copy_stream.synchronize()
