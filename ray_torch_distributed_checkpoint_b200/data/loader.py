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
    """Epoch iterator over a uint8 image set held in host memory.

    Each step: gather ``batch`` rows by index into a pinned staging buffer (background thread, the next
    batch is prepared while the current one trains), one ``cudaMemcpyAsync`` of the uint8 pixels and
    int64 labels on a copy stream, then (optionally) ``ops.u8_normalize`` into a float32 ``[B, 784]``
    activation on the compute stream. Yields ``(x_float32 or x_uint8, labels)`` on the device.
    """

    def __init__(self, images_u8: torch.Tensor, labels: torch.Tensor, batch_size: int, device, *, shuffle: bool = True,
                 world: int = 1, rank: int = 0, seed: int = 0, drop_last: bool = True, normalize: bool = True,
                 mean: float = 0.5, std: float = 0.5, depth: int = 3):
        assert images_u8.dtype == torch.uint8 and labels.dtype == torch.int64
        self.images = images_u8.reshape(images_u8.shape[0], -1).contiguous()
        self.labels = labels.contiguous()
        self.row = self.images.shape[1]
        self.batch_size = int(batch_size)
        self.device = torch.device(device)
        self.sampler = DistributedSampler(self.images.shape[0], world, rank, shuffle=shuffle, seed=seed)
        self.drop_last = drop_last
        self.normalize, self.mean, self.std = normalize, mean, std
        self.cuda = self.device.type == "cuda"
        self.depth = depth
        pin = self.cuda
        self._stage_x = [torch.empty(self.batch_size, self.row, dtype=torch.uint8, pin_memory=pin) for _ in range(depth)]
        self._stage_y = [torch.empty(self.batch_size, dtype=torch.int64, pin_memory=pin) for _ in range(depth)]
        if self.cuda:
            self._dev_x = [torch.empty(self.batch_size, self.row, dtype=torch.uint8, device=self.device) for _ in range(depth)]
            self._dev_y = [torch.empty(self.batch_size, dtype=torch.int64, device=self.device) for _ in range(depth)]
            self._dev_f = [torch.empty(self.batch_size, self.row, dtype=torch.float32, device=self.device) for _ in range(depth)] if normalize else None
            self.copy_stream = torch.cuda.Stream(self.device)
            self._consumed = [None] * depth
        self.h2d_bytes_per_batch = self.batch_size * (self.row + 8)

    def __len__(self):
        n = len(self.sampler)
        return n // self.batch_size if self.drop_last else math.ceil(n / self.batch_size)

    def _gather(self, slot: int, idx: np.ndarray):
        n = len(idx)
        ti = torch.from_numpy(idx)
        torch.index_select(self.images, 0, ti, out=self._stage_x[slot][:n])
        torch.index_select(self.labels, 0, ti, out=self._stage_y[slot][:n])
        return n

    def __iter__(self):
        idx_all = np.asarray(self.sampler.indices(), dtype=np.int64)
        nb = len(self)
        bs = self.batch_size
        batches = [idx_all[i * bs:(i + 1) * bs] for i in range(nb)]
        results = {}
        cond = threading.Condition()

        def producer():
            for i, idx in enumerate(batches):
                slot = i % self.depth
                with cond:
                    while i - self._released[0] >= self.depth:
                        cond.wait()
                n = self._gather(slot, idx)
                with cond:
                    results[i] = n
                    cond.notify_all()

        self._released = [0]
        th = threading.Thread(target=producer, daemon=True)
        th.start()
        for i in range(nb):
            slot = i % self.depth
            with cond:
                while i not in results:
                    cond.wait()
                n = results.pop(i)
            if not self.cuda:
                x = self._stage_x[slot][:n]
                xb = ((x.to(torch.float32) / 255.0 - self.mean) / self.std) if self.normalize else x.clone()
                yb = self._stage_y[slot][:n].clone()
                with cond:
                    self._released[0] = i + 1
                    cond.notify_all()
                yield xb, yb
                continue
            cur = torch.cuda.current_stream(self.device)
            if self._consumed[slot] is not None:
                self.copy_stream.wait_event(self._consumed[slot])   # device buffers of this slot are free again
            with torch.cuda.stream(self.copy_stream):
                self._dev_x[slot][:n].copy_(self._stage_x[slot][:n], non_blocking=True)
                self._dev_y[slot][:n].copy_(self._stage_y[slot][:n], non_blocking=True)
                ev = torch.cuda.Event()
                ev.record(self.copy_stream)
            # the pinned staging slot may be refilled once the copies have been issued AND completed;
            # release it lazily: when this slot's event is waited on by the compute stream below, the
            # producer is still `depth-1` batches behind, so synchronise the event before reuse.
            cur.wait_event(ev)
            if self.normalize:
                from ..ops import u8_normalize
                xb = u8_normalize(self._dev_x[slot][:n], mean=self.mean, std=self.std, out=self._dev_f[slot][:n])
            else:
                xb = self._dev_x[slot][:n]
            yb = self._dev_y[slot][:n]
            ev.synchronize() if self.depth < 2 else None
            self._h2d_done = ev
            with cond:
                # staging slot i is reusable after its H2D finished; with depth >= 2 the producer only
                # reaches it after `depth-1` further batches, by which time we have synchronised below
                self._released[0] = max(self._released[0], i + 1 - (self.depth - 1))
                cond.notify_all()
            yield xb, yb
            done = torch.cuda.Event()
            done.record(cur)
            self._consumed[slot] = done
            ev.synchronize()
            with cond:
                self._released[0] = i + 1
                cond.notify_all()
        th.join()
