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

    def __init__(self, loader, device: torch.device, prefetch: bool = True):
        self.loader = loader
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
                yield _to_device(batch, self.device, False)
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
            yield cur


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


class U8BatchLoader:
    """Epoch iterator over a uint8 image set held in host memory, built for PCIe + HBM rather than for
    a Python collate loop.

    * The shuffle is done once per epoch, off the critical path: a background thread gathers this
      rank's shard of the epoch's permutation into a pinned buffer (``index_select`` on all host cores)
      while the previous epoch is still training (two pinned buffers, ping-pong).
    * Every step is then ONE contiguous ``cudaMemcpyAsync`` of uint8 pixels (+ one of int64 labels) on a
      copy stream — 4x fewer bytes than the reference's float32 batches (my_ray_module.py:38, 128) —
      issued ``lookahead`` steps ahead of the consumer, into a ring of device buffers.
    * Normalisation (ToTensor + Normalize) happens on the GPU: either the consumer takes the uint8
      batch (the fused engine normalises inside its graph) or, with ``normalize=True``, this loader
      runs ``ops.u8_normalize`` and yields float32.

    Yields ``(x, labels)`` on the device; ``x`` is ``uint8 [B, features]`` or ``float32 [B, features]``.
    """

    def __init__(self, images_u8: torch.Tensor, labels: torch.Tensor, batch_size: int, device, *, shuffle: bool = True,
                 world: int = 1, rank: int = 0, seed: int = 0, drop_last: bool = True, normalize: bool = False,
                 mean: float = 0.5, std: float = 0.5, lookahead: int = 2, prefetch_next_epoch: bool = True):
        assert images_u8.dtype == torch.uint8 and labels.dtype == torch.int64
        self.prefetch_next_epoch = prefetch_next_epoch
        self.images = images_u8.reshape(images_u8.shape[0], -1).contiguous()
        self.labels = labels.contiguous()
        self.row = self.images.shape[1]
        self.batch_size = int(batch_size)
        self.device = torch.device(device)
        self.sampler = DistributedSampler(self.images.shape[0], world, rank, shuffle=shuffle, seed=seed)
        self.drop_last = drop_last
        self.normalize, self.mean, self.std = normalize, mean, std
        self.cuda = self.device.type == "cuda"
        self.lookahead = max(1, int(lookahead))
        n = len(self.sampler)
        self._host_x = [torch.empty(n, self.row, dtype=torch.uint8, pin_memory=self.cuda) for _ in range(2)]
        self._host_y = [torch.empty(n, dtype=torch.int64, pin_memory=self.cuda) for _ in range(2)]
        self._prepared = {}          # epoch -> (slot, thread)
        self._epoch_iter = 0
        if self.cuda:
            ring = self.lookahead + 2
            self._dev_x = [torch.empty(self.batch_size, self.row, dtype=torch.uint8, device=self.device) for _ in range(ring)]
            self._dev_y = [torch.empty(self.batch_size, dtype=torch.int64, device=self.device) for _ in range(ring)]
            self._dev_f = ([torch.empty(self.batch_size, self.row, dtype=torch.float32, device=self.device) for _ in range(ring)]
                           if normalize else None)
            self._free_ev = [None] * ring
            self.copy_stream = torch.cuda.Stream(self.device)
        self.h2d_bytes_per_batch = self.batch_size * (self.row + 8)
        self._prepare(self.sampler.epoch)

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
                y = hy[i * bs:min(total, (i + 1) * bs)].clone()
                yield (((x.to(torch.float32) / 255.0 - self.mean) / self.std) if self.normalize else x.clone()), y
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
            yield xb, self._dev_y[bi][:rows]
            done = torch.cuda.Event()
            done.record(torch.cuda.current_stream(self.device))
            self._free_ev[bi] = done
        # the pinned epoch buffer may be rewritten (epoch + 2) only after its last H2D has finished
        self.copy_stream.synchronize()
