"""Splitter-only timing: debug 3 (no TMA, no MMA) vs 19 (+ no proxy fence) vs 11 (no split at all)."""
# NOTE: needs a library built with the debug hooks: RTDC_EXTRA_NVCC_FLAGS=-DRTDC_GEMM_DEBUG_HOOKS python -m ray_torch_distributed_checkpoint_b200.ops._build --force

import os, sys
import torch
from ray_torch_distributed_checkpoint_b200 import ops
B = 18944
dev = "cuda"
x = torch.randint(0, 256, (B, 784), device=dev).float(); w1 = torch.randn(512, 784, device=dev) * .03; b1 = torch.randn(512, device=dev)
out = torch.empty(B, 512, device=dev)
def t(fn, n=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for _ in range(n): fn()
        g.replay(); s.synchronize()
        a.record(s); g.replay(); b.record(s); s.synchronize()
    return a.elapsed_time(b) / n * 1000
for ex in (False, True):
    row = []
    for dbg in [int(v) for v in (sys.argv[1:] or ["3", "19", "11", "7", "23", "15"])]:
        os.environ["RTDC_GEMM_DEBUG"] = str(dbg)
        us = t(lambda: ops.gemm(x, w1, out, bias=b1, relu=True, precision=3, a_exact=ex, bn=256))
        row.append(f"dbg{dbg}={us:.1f}")
    print(f"a_exact={int(ex)}: " + " ".join(row), flush=True)
