"""wgrad1 / wgrad2 timing (graph replays over rotating operands) for the split-K plan experiments."""
# NOTE: needs a library built with the debug hooks: RTDC_EXTRA_NVCC_FLAGS=-DRTDC_GEMM_DEBUG_HOOKS python -m ray_torch_distributed_checkpoint_b200.ops._build --force

import os, sys
import torch
from ray_torch_distributed_checkpoint_b200 import ops
B = 18944; dev = "cuda"; ROT = 4
xs = [torch.randint(0, 256, (B, 784), device=dev).float() - 127.5 for _ in range(ROT)]
hs = [torch.relu(torch.randn(B, 512, device=dev)) for _ in range(ROT)]
dys = [torch.randn(B, 512, device=dev) for _ in range(ROT)]
dw1 = torch.empty(512, 784, device=dev); dw2 = torch.empty(512, 512, device=dev)
def t(fns, reps=5):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for f in fns: f()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for f in fns: f()
        for _ in range(3): g.replay()
        s.synchronize()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(s); g.replay(); b.record(s); s.synchronize()
            ts.append(a.elapsed_time(b) / len(fns) * 1000)
    return sorted(ts)[len(ts) // 2]
for sl in (sys.argv[1:] or ["auto", "2", "3", "4", "5", "6", "7", "8", "9"]):
    if sl == "auto": os.environ.pop("RTDC_GEMM_SPLITS_LAST", None)
    else: os.environ["RTDC_GEMM_SPLITS_LAST"] = sl
    us = t([(lambda i=i: ops.linear_wgrad(dys[i], xs[i], out=dw1, x_exact=True, alpha=2 / 255)) for i in range(ROT)])
    ref = (2 / 255) * (dys[ROT - 1].double().T @ xs[ROT - 1].double())
    err = (dw1.double() - ref).abs().max().item() / ref.abs().max().item()
    print(f"wgrad1 exact splits_last={sl}: {us:.1f} us  relerr {err:.2e}", flush=True)
os.environ.pop("RTDC_GEMM_SPLITS_LAST", None)
print(f"wgrad2: {t([(lambda i=i: ops.linear_wgrad(dys[i], hs[i], out=dw2)) for i in range(ROT)]):.1f} us")
