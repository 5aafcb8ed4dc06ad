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

    def indices_tensor(self, dtype=torch.int64) -> torch.Tensor:
        """:meth:`indices` without Python lists (same values): what a loader over hundreds of thousands of rows wants."""
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            idx = torch.randperm(self.n, generator=g)
        else:
            idx = torch.arange(self.n)
        if not self.drop_last:
            pad = self.total_size - self.n
            if pad > 0:
                idx = torch.cat([idx] + [idx] * (pad // max(1, self.n)) + [idx[:pad % max(1, self.n)]])
        else:
            idx = idx[:self.total_size]
        return idx[self.rank:self.total_size:self.num_replicas].to(dtype).contiguous()

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


def _sync_ranks_for_stamp():
    """Before the first timed batch of an epoch: wait until every rank is here. Ranks leave an epoch boundary at different
    times (rank 0 snapshots the checkpoint and does the bookkeeping); without this the other ranks' first step of the epoch
    holds that wait inside its gradient-reduction kernel and it would be counted as device time of the step. The benchmark's
    reference arm gets the same treatment on every step (baseline/ray_standin)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.barrier()


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


class BlockSampler:
    """Two-level shuffle for the streaming transfer: the epoch's order is a permutation of BLOCKS of ``block`` consecutive
    rows (of a dataset that was row-shuffled once when the loader was built), sharded over ranks like
    ``DistributedSampler`` shards rows: the block list is padded to a multiple of the world size by wrapping around and rank
    r takes every world-th block. Every block has exactly ``block`` rows (the last one is ``[N - block, N)``, i.e. it
    overlaps its neighbour instead of being short), so all ranks see the same number of rows. A batch is then a run of
    contiguous row ranges — something copy engines move without any host-side gather."""

    def __init__(self, dataset_len: int, block: int, num_replicas: int, rank: int, shuffle: bool = True, seed: int = 0):
        self.n = int(dataset_len)
        self.block = max(1, min(int(block), self.n))
        self.num_replicas, self.rank, self.shuffle, self.seed, self.epoch = num_replicas, rank, shuffle, seed, 0
        self.nblocks = math.ceil(self.n / self.block)
        self.blocks_per_rank = math.ceil(self.nblocks / num_replicas)

    def set_epoch(self, epoch: int) -> None:
        self.epoch = int(epoch)

    def block_starts(self) -> list:
        """First row of each block of this rank's shard for the current epoch."""
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            order = torch.randperm(self.nblocks, generator=g).tolist()
        else:
            order = list(range(self.nblocks))
        total = self.blocks_per_rank * self.num_replicas
        pad = total - len(order)
        if pad > 0:
            order += (order * math.ceil(pad / len(order)))[:pad]
        mine = order[self.rank:total:self.num_replicas]
        last = self.n - self.block
        return [min(b * self.block, last) for b in mine]

    def __len__(self) -> int:
        return self.blocks_per_rank * self.block

    def __iter__(self):
        for s in self.block_starts():
            yield from range(s, s + self.block)


def shared_pinned_rows(rows: torch.Tensor, order: Optional[torch.Tensor], world: int, rank: int) -> torch.Tensor:
    """``rows[order]`` (or ``rows``) as ONE page-locked host buffer per node, mapped by every rank, instead of one private
    pinned copy per rank: the node's first rank writes it into a POSIX shared-memory file, every rank maps that file and
    registers the mapping with the CUDA driver (``cudaHostRegister``), after which copy engines of all GPUs read the same
    physical pages. Collective over the default process group when ``world > 1``; a private pinned tensor otherwise."""
    import torch.distributed as dist
    n, row = rows.shape
    if world <= 1 or not (dist.is_available() and dist.is_initialized() and dist.get_world_size() == world):
        out = torch.empty(n, row, dtype=rows.dtype, pin_memory=True)
        if order is not None:
            torch.index_select(rows, 0, order, out=out)
        else:
            out.copy_(rows)
        return out
    import os
    import uuid
    box = [f"/dev/shm/rtdc_ds_{os.getpid()}_{uuid.uuid4().hex[:8]}" if rank == 0 else None]
    dist.broadcast_object_list(box, src=0)
    path, nbytes = box[0], n * row * rows.element_size()
    if rank == 0:
        flat = torch.from_file(path, shared=True, size=nbytes, dtype=torch.uint8)
        view = flat.view(rows.dtype).view(n, row)
        if order is not None:
            torch.index_select(rows, 0, order, out=view)
        else:
            view.copy_(rows)
    dist.barrier()
    if rank != 0:
        flat = torch.from_file(path, shared=True, size=nbytes, dtype=torch.uint8)
        view = flat.view(rows.dtype).view(n, row)
    rc = torch.cuda.cudart().cudaHostRegister(flat.data_ptr(), nbytes, 0)
    if int(rc) != 0:
        raise RuntimeError(f"cudaHostRegister of the shared dataset failed with code {int(rc)}")
    dist.barrier()
    if rank == 0:
        os.unlink(path)                      # the mappings keep the pages alive
    return view


_live_loaders = None


def _register_for_exit(loader):
    """Background shuffle threads must not be inside torch's thread pool when the interpreter tears down."""
    global _live_loaders
    if _live_loaders is None:
        import atexit
        import weakref
        _live_loaders = weakref.WeakSet()

        def _join_all():
            for ld in list(_live_loaders):
                ld.close()
        atexit.register(_join_all)
    _live_loaders.add(loader)


class U8BatchLoader:
    """Epoch iterator over a uint8 image set, built for PCIe + HBM rather than for a Python collate loop. Pixels and
    labels travel as bytes (4x fewer pixel bytes than the reference's float32 batches, 8x fewer label bytes than int64;
    my_ray_module.py:38,128), batches are assembled ``lookahead`` steps ahead of the consumer on a side stream into a ring
    of device buffers, and normalisation (ToTensor + Normalize) happens on the GPU. Three ways to get a shuffled batch onto
    the device (``transfer``):

    * ``"resident"`` — the dataset is uploaded ONCE and stays in HBM (all of FashionMNIST is 47 MB of a 180 GB device).
      Per epoch the sampler's permutation (``DistributedSampler`` arithmetic, exact row-level shuffle) is uploaded as
      int32 indices; per step ``ops.u8_gather_rows`` assembles the batch at HBM speed. ``"auto"`` picks this whenever the
      dataset fits ``resident_max_bytes``.
    * ``"stream"`` — the dataset stays in pinned host memory and every step's pixels cross PCIe by DMA. To need no
      host-side gather the shuffle has two levels (:class:`BlockSampler`): rows are shuffled once when the loader is
      built, and each epoch permutes and shards blocks of ``shuffle_block`` consecutive rows; a batch is ~B/shuffle_block
      contiguous pieces, one ``cudaMemcpyAsync`` each (``ops.h2d_pieces``). The block size trades shuffle granularity for
      copy-engine efficiency (measured on a B200, 14.9 MB per step, ``tools/input_probe.py``): 512-row pieces (0.4 MB) move
      33 GB/s, 2048-row pieces (1.6 MB, the default) 47 GB/s, one contiguous copy 54 GB/s — a 0.33 ms training step needs 45.
    * ``"host_gather"`` — exact row-level shuffle with the dataset on the host: a background thread gathers the rank's
      shard of the epoch's permutation into a pinned buffer (``index_select``; ping-pong with the next epoch), each step is
      one contiguous copy. Exact and simple, but a host gathers 784-byte rows at a few GB/s: at B200 step times it
      is the slowest of the three by far.

    Yields ``(x, labels)`` on the device; ``x`` is ``uint8 [B, features]``, :class:`U8Images` ``[B, *image_shape]``
    (``wrap=True``) or ``float32 [B, features]`` (``normalize=True``); ``labels`` is ``int64 [B]``.
    """

    def __init__(self, images_u8: torch.Tensor, labels: torch.Tensor, batch_size: int, device, *, shuffle: bool = True,
                 world: int = 1, rank: int = 0, seed: int = 0, drop_last: bool = True, normalize: bool = False,
                 mean: float = 0.5, std: float = 0.5, lookahead: int = 2, prefetch_next_epoch: bool = True,
                 wrap: bool = False, timed: bool = False, transfer: str = "auto", shuffle_block: int = 2048,
                 resident_max_bytes: int = 4 << 30, resident_host_perm_max_rows: int = 250_000):
        assert images_u8.dtype == torch.uint8 and labels.dtype in (torch.int64, torch.uint8)
        self.prefetch_next_epoch = prefetch_next_epoch
        self.image_shape = tuple(images_u8.shape[1:]) if images_u8.dim() > 2 else (images_u8.shape[1],)
        if len(self.image_shape) == 2:
            self.image_shape = (1,) + self.image_shape          # ToTensor's channel dimension
        images = images_u8.reshape(images_u8.shape[0], -1)
        # labels travel as bytes when they fit (10 classes do)
        self.label_dtype = torch.uint8 if (labels.dtype == torch.uint8 or (labels.numel() and 0 <= int(labels.min()) and int(labels.max()) < 256)) else torch.int64
        labels = labels.to(self.label_dtype)
        self.row = images.shape[1]
        self.batch_size = int(batch_size)
        self.device = torch.device(device)
        self.cuda = self.device.type == "cuda"
        n_rows = images.shape[0]
        if transfer not in ("auto", "resident", "stream", "host_gather"):
            raise ValueError(f"transfer must be 'auto', 'resident', 'stream' or 'host_gather' (got {transfer!r})")
        if not self.cuda:
            transfer = "host_gather"
        elif transfer in ("auto", "resident"):
            fits = n_rows * (self.row + 1) <= resident_max_bytes and self.row % 16 == 0 and self.label_dtype == torch.uint8
            if transfer == "resident" and not fits:
                raise ValueError("transfer='resident' needs uint8 labels, a row length that is a multiple of 16 bytes and a dataset within resident_max_bytes")
            transfer = "resident" if fits else "stream"
        self.transfer = transfer
        self.drop_last = drop_last
        self.normalize, self.mean, self.std = normalize, mean, std
        self.wrap, self.timed = wrap, timed
        self.lookahead = max(1, int(lookahead))
        self._prepared = {}          # epoch -> (slot, thread, result box) of the background plan
        self._closed = False
        self._iterated_epoch = None
        self._epoch_copies_done = [None, None]
        self._device_perm = False

        if transfer == "stream":
            self.sampler = BlockSampler(n_rows, shuffle_block, world, rank, shuffle=shuffle, seed=seed)
            if shuffle:                                      # level one: rows, once
                g = torch.Generator()
                g.manual_seed(seed + 0x5EED)
                order = torch.randperm(n_rows, generator=g)
                self.images = shared_pinned_rows(images, order, world, rank)
                self.labels = torch.index_select(labels, 0, order).contiguous()
            else:
                self.images = images.contiguous() if images.is_pinned() else shared_pinned_rows(images, None, world, rank)
                self.labels = labels.contiguous()
            n = len(self.sampler)
            self._host_y = [torch.empty(n, dtype=self.label_dtype, pin_memory=True) for _ in range(2)]
            self.h2d_bytes_per_batch = self.batch_size * (self.row + (1 if self.label_dtype == torch.uint8 else 8))
        else:
            self.sampler = DistributedSampler(n_rows, world, rank, shuffle=shuffle, seed=seed)
            n = len(self.sampler)
            if transfer == "resident":
                self.images = images.contiguous().to(self.device)            # the one upload of the pixels
                self.labels = labels.contiguous().to(self.device)
                # torch.randperm on one host core costs ~12 ns per row: fine for FashionMNIST (0.7 ms), not for millions of rows
                # per epoch, where the permutation is drawn on the GPU instead
                self._device_perm = shuffle and n_rows > resident_host_perm_max_rows
                self._host_idx = [torch.empty(n, dtype=torch.int32, pin_memory=True) for _ in range(2)]
                self._dev_idx = [torch.empty(n, dtype=torch.int32, device=self.device) for _ in range(2)]
                self.h2d_bytes_per_batch = self.batch_size * 4                # the batch's indices
            else:
                self.images = images.contiguous()
                self.labels = labels.contiguous()
                self._host_x = [torch.empty(n, self.row, dtype=torch.uint8, pin_memory=self.cuda) for _ in range(2)]
                self._host_y = [torch.empty(n, dtype=self.label_dtype, pin_memory=self.cuda) for _ in range(2)]
                self.h2d_bytes_per_batch = self.batch_size * (self.row + (1 if self.label_dtype == torch.uint8 else 8))
        if self.cuda:
            ring = self.lookahead + 2
            self._dev_x = [torch.empty(self.batch_size, self.row, dtype=torch.uint8, device=self.device) for _ in range(ring)]
            self._dev_y = [torch.empty(self.batch_size, dtype=torch.int64, device=self.device) for _ in range(ring)]
            self._dev_y8 = ([torch.empty(self.batch_size, dtype=torch.uint8, device=self.device) for _ in range(ring)]
                            if (self.label_dtype == torch.uint8 and transfer != "resident") else None)
            self._dev_f = ([torch.empty(self.batch_size, self.row, dtype=torch.float32, device=self.device) for _ in range(ring)]
                           if normalize else None)
            self._free_ev = [None] * ring
            self.copy_stream = torch.cuda.Stream(self.device)
        if timed:
            TIMING["extra"]["h2d_bytes_per_step"] = self.h2d_bytes_per_batch
            TIMING["extra"]["input_transfer"] = transfer
        _register_for_exit(self)
        self._prepare(self.sampler.epoch)

    @property
    def dataset(self):
        return self.images

    def __len__(self):
        n = len(self.sampler)
        return n // self.batch_size if self.drop_last else math.ceil(n / self.batch_size)

    def close(self):
        """Join the background epoch-plan threads."""
        self._closed = True
        for e in list(self._prepared):
            self._prepared.pop(e)[1].join()

    # ---------------------------------------------------------------- per-epoch host-side plan (background thread)
    def _plan(self, epoch: int, slot: int):
        """Everything the host has to compute for an epoch, done off the training thread while the previous epoch runs:
        the permutation (resident: int32 indices into a pinned buffer; stream: block order + the shard's labels;
        host_gather: the gathered pixel / label shard)."""
        old_epoch = self.sampler.epoch
        self.sampler.set_epoch(epoch)
        try:
            if self.transfer == "stream":
                starts = torch.as_tensor(self.sampler.block_starts(), dtype=torch.int64)
                G = self.sampler.block
                torch.index_select(self.labels.view(-1), 0, (starts[:, None] + torch.arange(G)[None, :]).reshape(-1), out=self._host_y[slot])
                return starts
            if self.transfer == "resident" and self._device_perm:
                return "device"                     # the permutation itself is drawn on the GPU in _epoch_resident
            idx = self.sampler.indices_tensor()
        finally:
            self.sampler.set_epoch(old_epoch)
        if self.transfer == "resident":
            self._host_idx[slot].copy_(idx)
            return None
        torch.index_select(self.images, 0, idx, out=self._host_x[slot])
        torch.index_select(self.labels, 0, idx, out=self._host_y[slot])
        return None

    def _prepare(self, epoch: int):
        if epoch in self._prepared or self._closed:
            return
        slot = epoch % 2
        box = {}

        done = self._epoch_copies_done[slot] if self.cuda else None

        def work():
            try:
                if done is not None:
                    done.synchronize()          # the H2D copies of epoch - 2 out of this slot's pinned buffers
                box["plan"] = self._plan(epoch, slot)
            except BaseException as e:          # surfaced when the epoch is iterated
                box["error"] = e

        th = threading.Thread(target=work, name="rtdc-epoch-plan", daemon=True)
        th.start()
        self._prepared[epoch] = (slot, th, box)

    def _take_plan(self, epoch: int):
        """Join (or compute) the plan of ``epoch``; start the next epoch's."""
        self._prepare(epoch)
        slot, th, box = self._prepared.pop(epoch)
        th.join()
        for e in list(self._prepared):          # stale preparations (epoch set backwards)
            if e != epoch + 1:
                self._prepared.pop(e)[1].join()
        if "error" in box:
            raise box["error"]
        if len(self) > 1 and self.prefetch_next_epoch and self.sampler.shuffle:
            self._prepare(epoch + 1)            # next epoch's shuffle overlaps this epoch's training
        return slot, box.get("plan")

    def _emit(self, xb: torch.Tensor, yb: torch.Tensor, ready=None):
        if self.normalize or not self.wrap:
            return xb, yb
        x = U8Images.wrap(xb.view(xb.shape[0], *self.image_shape), self.mean, self.std)
        x._ready = ready              # the event after which these bytes (and the labels) are in place
        return x, yb

    # ---------------------------------------------------------------- per-epoch set-up of each transfer
    def _epoch_host_gather(self, epoch: int):
        slot, _ = self._take_plan(epoch)
        hx, hy = self._host_x[slot], self._host_y[slot]

        def fill(bi, lo, hi):
            self._dev_x[bi][:hi - lo].copy_(hx[lo:hi], non_blocking=True)
            self._copy_labels(bi, hy, lo, hi)
        return fill, (hx, hy)

    def _copy_labels(self, bi, hy, lo, hi):
        if self._dev_y8 is not None:
            self._dev_y8[bi][:hi - lo].copy_(hy[lo:hi], non_blocking=True)
            self._dev_y[bi][:hi - lo].copy_(self._dev_y8[bi][:hi - lo])        # widen on the device
        else:
            self._dev_y[bi][:hi - lo].copy_(hy[lo:hi], non_blocking=True)

    def _device_indices(self, epoch: int, slot: int):
        """This rank's shard of the epoch's permutation, drawn on the GPU (same seed on every rank => same permutation; same
        pad-by-wrapping and stride-by-rank arithmetic as DistributedSampler, but not torch's CPU random stream)."""
        sm = self.sampler
        g = torch.Generator(device=self.device)
        g.manual_seed(sm.seed + epoch)
        perm = torch.randperm(sm.n, device=self.device, generator=g) if sm.shuffle else torch.arange(sm.n, device=self.device)
        pad = sm.total_size - sm.n
        if pad > 0:
            perm = torch.cat([perm] + [perm] * (pad // sm.n) + [perm[:pad % sm.n]])
        self._dev_idx[slot].copy_(perm[sm.rank:sm.total_size:sm.num_replicas])

    def _epoch_resident(self, epoch: int):
        from ..ops import u8_gather_rows
        slot, plan = self._take_plan(epoch)
        with torch.cuda.stream(self.copy_stream):
            if plan == "device":
                self._device_indices(epoch, slot)
            else:
                self._dev_idx[slot].copy_(self._host_idx[slot], non_blocking=True)
        idx = self._dev_idx[slot]

        def fill(bi, lo, hi):
            u8_gather_rows(self.images, idx[lo:hi], self._dev_x[bi], labels=self.labels, y_out=self._dev_y[bi], stream=self.copy_stream)
        return fill, None

    def _epoch_stream(self, epoch: int):
        from ..ops import h2d_pieces
        slot, starts = self._take_plan(epoch)
        G = self.sampler.block
        hy = self._host_y[slot]
        row = self.row
        starts_l = starts.tolist()

        def fill(bi, lo, hi):
            b0, b1 = lo // G, (hi - 1) // G
            offs = [s_ * row for s_ in starts_l[b0:b1 + 1]]
            sizes = [G * row] * (b1 - b0 + 1)
            head = lo - b0 * G                                   # the window may start / end inside a block
            offs[0] += head * row
            sizes[0] -= head * row
            sizes[-1] -= ((b1 + 1) * G - hi) * row
            h2d_pieces(self._dev_x[bi], self.images, torch.tensor(offs, dtype=torch.int64), torch.tensor(sizes, dtype=torch.int64),
                       stream=self.copy_stream)
            self._copy_labels(bi, hy, lo, hi)
        return fill, None

    def __iter__(self):
        epoch = self.sampler.epoch
        if self.sampler.shuffle and self._iterated_epoch == epoch:
            # nobody called sampler.set_epoch since the last pass (single-worker loops do not, my_ray_module.py:149-151):
            # reshuffle anyway, like the RandomSampler of the DataLoader this loader stands in for
            epoch += 1
            self.sampler.set_epoch(epoch)
        self._iterated_epoch = epoch
        nb, bs = len(self), self.batch_size
        total = len(self.sampler)
        if not self.cuda:
            _, (hx, hy) = self._epoch_host_gather(epoch)
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
        fill, _ = {"host_gather": self._epoch_host_gather, "resident": self._epoch_resident, "stream": self._epoch_stream}[self.transfer](epoch)
        ring = len(self._dev_x)
        inflight = []                            # (buffer index, rows, batch-ready event)
        issued = 0

        def issue():
            nonlocal issued
            i = issued
            bi = i % ring
            lo, hi = i * bs, min(total, (i + 1) * bs)
            if self._free_ev[bi] is not None:
                self.copy_stream.wait_event(self._free_ev[bi])      # consumer is done with this device buffer
            with torch.cuda.stream(self.copy_stream):
                fill(bi, lo, hi)
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
                if _ == 0:
                    _sync_ranks_for_stamp()
                _stamp("step_begin", self.device)
            yield self._emit(xb, self._dev_y[bi][:rows], ready=ev)
            if self.timed:
                _stamp("step_end", self.device)
            done = torch.cuda.Event()
            done.record(torch.cuda.current_stream(self.device))
            self._free_ev[bi] = done
        # pinned per-epoch buffers may be rewritten (by the plan of epoch + 2) only after their last H2D has finished: that plan
        # thread waits for this event, the training thread does not (it goes on to validation while the GPU drains the epoch)
        ev = torch.cuda.Event()
        ev.record(self.copy_stream)
        self._epoch_copies_done[epoch % 2] = ev
