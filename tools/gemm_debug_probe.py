"""Component timing of the GEMM kernel with RTDC_GEMM_DEBUG bits (1 no TMA, 2 no MMA, 4 no stores, 8 no split, 16 no fence)."""
# NOTE: needs a library built with the debug hooks: RTDC_EXTRA_NVCC_FLAGS=-DRTDC_GEMM_DEBUG_HOOKS python -m ray_torch_distributed_checkpoint_b200.ops._build --force

import os, sys
import torch
from ray_torch_distributed_checkpoint_b200 import ops
B = 18944
dev = "cuda"
x = torch.randn(B, 784, device=dev); w1 = torch.randn(512, 784, device=dev) * .03; b1 = torch.randn(512, device=dev)
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
for dbg in [int(v) for v in (sys.argv[1:] or ['0', '8', '2', '10', '1', '3', '11'])]:
    os.environ["RTDC_GEMM_DEBUG"] = str(dbg)
    us = t(lambda: ops.gemm(x, w1, out, bias=b1, relu=True, precision=3))
    print(f"fwd1 prec=3 B={B} debug={dbg}: {us:.1f} us", flush=True)
