"""Whole-step CUDA graphs for models outside the MLP engine (the GPT-2 / ResNet-50 configurations of BASELINE.json).

The reference MLP gets its single-graph step from ``parallel.MLPTrainEngine``; a general module trained through autograd
launches a few hundred kernels per step from Python, and on a B200 the GPU finishes them faster than the host can queue
them for stretches of the step. ``GraphedStep`` captures one complete step — gradient reset, forward, loss, backward,
optimizer — into one CUDA graph and replays it: no tracing compiler, the same kernels in the same order, launched by the
driver instead of by Python. Everything this framework launches is capturable by construction (kernels enqueue on the
current stream, tensor maps travel as kernel parameters, optimizer step counters live on the device).
"""
from __future__ import annotations

from typing import Callable, Optional

import torch


class GraphedStep:
    """``step = GraphedStep(fn, warmup=3)``; ``out = step()`` replays the captured graph of ``fn()``.

    ``fn`` must be *static*: it reads its inputs from tensors that live across calls (refresh them with ``copy_`` before a
    call), takes no data-dependent host decisions, and does not synchronise with the host. ``fn``'s return value (e.g. the
    loss tensor) is captured once and overwritten by every replay. Gradients must stay allocated between steps
    (``zero_grad(set_to_none=False)``; this framework's optimizers keep them as views of one arena). On a CPU-only build, or
    with ``enabled=False``, calls run ``fn`` eagerly.
    """

    def __init__(self, fn: Callable[[], object], warmup: int = 3, enabled: bool = True, device: Optional[torch.device] = None):
        self.fn = fn
        self.graph = None
        self.out = None
        self.enabled = enabled and torch.cuda.is_available()
        if not self.enabled:
            return
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):                  # lazy initialisations (library handles, autotuning, allocator growth)
            for _ in range(max(1, warmup)):
                fn()
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.out = fn()

    def __call__(self):
        if self.graph is None:
            return self.fn()
        self.graph.replay()
        return self.out
