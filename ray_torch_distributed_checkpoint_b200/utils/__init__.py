"""Timing, tracing and GPU sampling helpers: CUDA-event step timers with max-over-ranks,
NVTX ranges, an NVML sampler (the ``@gpu_profile`` equivalent); whole-step CUDA graphs for general modules."""
from .graphs import GraphedStep
from .profiling import GpuSampler, StepTimer, max_over_ranks, nvtx_range

__all__ = ["GraphedStep", "GpuSampler", "StepTimer", "max_over_ranks", "nvtx_range"]
