"""RTDC_GEMM_DEBUG matrix for the wgrad2 shape (both operands MN-major, split-K) next to the fwd2 shape (K-major)."""
# NOTE: needs a library built with the debug hooks: RTDC_EXTRA_NVCC_FLAGS=-DRTDC_GEMM_DEBUG_HOOKS python -m ray_torch_distributed_checkpoint_b200.ops._build --force

import os, sys
import torch
from ray_torch_distributed_checkpoint_b200 import ops
B = 18944; dev = "cuda"
h = torch.relu(torch.randn(B, 512, device=dev)); dy = torch.randn(B, 512, device=dev); w2 = torch.randn(512, 512, device=dev) * .03
dw2 = torch.empty(512, 512, device=dev); out = torch.empty(B, 512, device=dev)
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
cases = {"wgrad2": lambda: ops.linear_wgrad(dy, h, out=dw2), "fwd2": lambda: ops.gemm(h, w2, out, relu=True),
         "dgrad2": lambda: ops.linear_dgrad(dy, w2, out=out)}
for name, fn in cases.items():
    row = []
    for dbg in (0, 8, 2, 10, 9, 3, 11):
        os.environ["RTDC_GEMM_DEBUG"] = str(dbg)
        row.append(f"dbg{dbg}={t(fn):.1f}")
    print(name, " ".join(row), flush=True)
