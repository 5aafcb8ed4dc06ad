"""ray.train.torch stand-in: stock DDP + NCCL, DistributedSampler, a device-moving loader wrapper that
doubles as the benchmark's stopwatch (CUDA events at train-batch boundaries; see baseline/README.md)."""
from __future__ import absolute_import

import os
import time

import torch as _torch
import torch.distributed as _dist
from torch.utils.data import DataLoader, DistributedSampler, RandomSampler

from . import _state, Result

TIMING = {"records": [], "loaders": 0}     # filled by the first (training) loader


def get_device():
    if _state["use_gpu"] and _torch.cuda.is_available():
        return _torch.device("cuda", int(os.environ.get("LOCAL_RANK", 0)))
    return _torch.device("cpu")


def prepare_model(model, **_):
    dev = get_device()
    model = model.to(dev)
    if _dist.is_initialized() and _dist.get_world_size() > 1:
        kw = {"device_ids": [dev.index]} if dev.type == "cuda" else {}
        model = _torch.nn.parallel.DistributedDataParallel(model, **kw)
    return model


class _DeviceLoader:
    def __init__(self, loader, device, timed):
        self.loader, self.device, self.timed = loader, device, timed

    @property
    def sampler(self):
        return self.loader.sampler

    def __len__(self):
        return len(self.loader)

    def __iter__(self):
        cuda = self.device.type == "cuda"
        rec = TIMING["records"] if self.timed else None
        for batch in self.loader:
            moved = tuple(t.to(self.device, non_blocking=True) for t in batch)
            if rec is not None:
                ev = _torch.cuda.Event(enable_timing=True) if cuda else None
                if cuda:
                    ev.record()
                rec.append(("step_begin", time.perf_counter(), ev))
                if cuda and _dist.is_initialized() and _dist.get_world_size() > 1:
                    # second begin stamp once EVERY rank holds its batch: without it a step's device interval at N >= 2
                    # also holds the slowest peer's CPU-loader stall (it surfaces inside the NCCL all-reduce)
                    _dist.barrier()
                    ev2 = _torch.cuda.Event(enable_timing=True)
                    ev2.record()
                    rec.append(("step_begin_synced", time.perf_counter(), ev2))
                else:
                    rec.append(("step_begin_synced", time.perf_counter(), ev))
            yield moved
            if rec is not None:
                ev = _torch.cuda.Event(enable_timing=True) if cuda else None
                if cuda:
                    ev.record()
                rec.append(("step_end", time.perf_counter(), ev))


def prepare_data_loader(data_loader, **_):
    world = _dist.get_world_size() if _dist.is_initialized() else 1
    if world > 1:
        shuffle = isinstance(data_loader.sampler, RandomSampler)
        sampler = DistributedSampler(data_loader.dataset, shuffle=shuffle)
        data_loader = DataLoader(data_loader.dataset, batch_size=data_loader.batch_size, sampler=sampler,
                                 num_workers=data_loader.num_workers, collate_fn=data_loader.collate_fn,
                                 pin_memory=data_loader.pin_memory, drop_last=data_loader.drop_last)
    timed = TIMING["loaders"] == 0          # the reference prepares the training loader first
    TIMING["loaders"] += 1
    return _DeviceLoader(data_loader, get_device(), timed)


class TorchTrainer:
    def __init__(self, train_loop_per_worker, train_loop_config=None, scaling_config=None, run_config=None, **_):
        self.fn, self.cfg, self.sc, self.rc = train_loop_per_worker, train_loop_config, scaling_config, run_config

    def fit(self):
        world = int(os.environ.get("WORLD_SIZE", 1))
        if world != self.sc.num_workers:
            raise ValueError(f"launch has WORLD_SIZE={world} but ScalingConfig.num_workers={self.sc.num_workers}")
        _state["use_gpu"] = bool(self.sc.use_gpu)
        if self.sc.use_gpu:
            _torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", 0)))
        if world > 1 and not _dist.is_initialized():
            _dist.init_process_group(backend="nccl" if self.sc.use_gpu else "gloo")
        run_dir = os.path.join(self.rc.storage_path, self.rc.name or "TorchTrainer_standin", "trial_0")
        os.makedirs(run_dir, exist_ok=True)
        _state.update(run_dir=run_dir, index=0, kept=[], num_to_keep=self.rc.checkpoint_config.num_to_keep,
                      last_metrics=None, last_ckpt=None)
        self.fn(self.cfg) if self.cfg is not None else self.fn()
        return Result(_state["last_metrics"], _state["last_ckpt"], run_dir)
