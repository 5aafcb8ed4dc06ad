"""Measurement helpers.

The reference times epochs with ``time.time()`` and prints minutes (``/root/reference/my_ray_module.py:144,
147,207-213``) and samples GPU utilisation with Metaflow's ``@gpu_profile`` (``train_flow.py:51``). Here:
device-side timing with CUDA events (the only clock that is valid for asynchronous kernels), reduced with
max over ranks as BASELINE.json requires; NVTX ranges for external profilers; an NVML sampler thread.
"""
from __future__ import annotations

import contextlib
import json
import threading
import time
from typing import Dict, List, Optional

import torch


def max_over_ranks(value: float, device=None, group=None) -> float:
    import torch.distributed as dist
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return float(value)
    dev = device if device is not None else (torch.device("cuda", torch.cuda.current_device()) if torch.cuda.is_available() else "cpu")
    t = torch.tensor([float(value)], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


class StepTimer:
    """Brackets regions of the current CUDA stream with events; ``summary()`` synchronises once and returns
    per-name totals in milliseconds (max over ranks when a process group exists)."""

    def __init__(self, enabled: bool = True):
        self.enabled = enabled and torch.cuda.is_available()
        self._open: Dict[str, torch.cuda.Event] = {}
        self._pairs: Dict[str, List] = {}
        self._cpu: Dict[str, float] = {}

    @contextlib.contextmanager
    def region(self, name: str):
        if not self.enabled:
            t0 = time.perf_counter()
            yield
            self._cpu[name] = self._cpu.get(name, 0.0) + (time.perf_counter() - t0) * 1e3
            return
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record()
        try:
            yield
        finally:
            b.record()
            self._pairs.setdefault(name, []).append((a, b))

    def summary(self, reduce_max: bool = True) -> Dict[str, float]:
        out = dict(self._cpu)
        if self.enabled:
            torch.cuda.synchronize()
            for name, pairs in self._pairs.items():
                out[name] = out.get(name, 0.0) + sum(a.elapsed_time(b) for a, b in pairs)
        if reduce_max:
            out = {k: max_over_ranks(v) for k, v in sorted(out.items())}
        return out

    def reset(self):
        self._pairs.clear()
        self._cpu.clear()


@contextlib.contextmanager
def nvtx_range(name: str):
    """NVTX range visible to external profilers; a no-op without CUDA."""
    if torch.cuda.is_available():
        torch.cuda.nvtx.range_push(name)
        try:
            yield
        finally:
            torch.cuda.nvtx.range_pop()
    else:
        yield


class GpuSampler:
    """Background NVML sampler: utilisation, memory, SM clock, power per GPU every ``interval`` seconds."""

    def __init__(self, interval: float = 1.0, path: Optional[str] = None):
        self.interval, self.path = max(0.02, float(interval)), path
        self.samples: List[Dict[str, float]] = []
        self._stop = threading.Event()
        self._thread = threading.Thread(target=self._run, name="rtdc-gpu-sampler", daemon=True)

    def _run(self):
        try:
            import pynvml
            pynvml.nvmlInit()
            handles = [pynvml.nvmlDeviceGetHandleByIndex(i) for i in range(pynvml.nvmlDeviceGetCount())]
        except Exception:
            return
        t0 = time.time()
        while not self._stop.is_set():
            row = {"t": round(time.time() - t0, 3)}
            for i, h in enumerate(handles):
                try:
                    u = pynvml.nvmlDeviceGetUtilizationRates(h)
                    row[f"gpu{i}_util"] = u.gpu
                    row[f"gpu{i}_mem_mib"] = pynvml.nvmlDeviceGetMemoryInfo(h).used // (1 << 20)
                    row[f"gpu{i}_sm_mhz"] = pynvml.nvmlDeviceGetClockInfo(h, pynvml.NVML_CLOCK_SM)
                    row[f"gpu{i}_power_w"] = pynvml.nvmlDeviceGetPowerUsage(h) / 1000.0
                except Exception:
                    pass
            self.samples.append(row)
            self._stop.wait(self.interval)

    def __enter__(self):
        self._thread.start()
        return self

    def __exit__(self, *exc):
        self._stop.set()
        self._thread.join(timeout=2.0)
        if self.path:
            with open(self.path, "w") as f:
                json.dump(self.samples, f)
