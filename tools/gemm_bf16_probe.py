"""bf16 GEMM on GPT-2 small's Linear shapes (16 x 1024 tokens): this repo's tcgen05 kernel vs torch (cuBLAS), forward / dgrad /
wgrad, device-timed with CUDA events around a CUDA graph of 20 calls over rotating operands. One GPU."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from ray_torch_distributed_checkpoint_b200 import ops  # noqa: E402

dev = "cuda"
M = 16384
REPS, ROT = 20, 4


def timed(fn):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(3):
            fn(i % ROT)
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for i in range(REPS):
                fn(i % ROT)
        g.replay(); s.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        a.record(s); g.replay(); b.record(s); s.synchronize()
    return a.elapsed_time(b) / REPS * 1e3


out = {}
for name, (cin, cout) in {"c_attn 768->2304": (768, 2304), "attn.c_proj 768->768": (768, 768), "c_fc 768->3072": (768, 3072),
                          "mlp.c_proj 3072->768": (3072, 768)}.items():
    xs = [torch.randn(M, cin, device=dev).to(torch.bfloat16) for _ in range(ROT)]
    dys = [torch.randn(M, cout, device=dev).to(torch.bfloat16) for _ in range(ROT)]
    w = (torch.randn(cout, cin, device=dev) * 0.02).to(torch.bfloat16)
    b = torch.zeros(cout, device=dev, dtype=torch.bfloat16)
    y = torch.empty(M, cout, device=dev, dtype=torch.bfloat16)
    dx = torch.empty(M, cin, device=dev, dtype=torch.bfloat16)
    dw32 = torch.empty(cout, cin, device=dev, dtype=torch.float32)
    dw16 = torch.empty(cout, cin, device=dev, dtype=torch.bfloat16)
    fl = 2.0 * M * cin * cout
    row = {}
    for op, ours, ref in (
            ("fwd", lambda i: ops.gemm_bf16(xs[i], w, out=y, bias=b), lambda i: torch.addmm(b, xs[i], w.t(), out=y)),
            ("dgrad", lambda i: ops.gemm_bf16(dys[i], w, out=dx, b_mn=True), lambda i: torch.mm(dys[i], w, out=dx)),
            ("wgrad", lambda i: ops.gemm_bf16(dys[i], xs[i], out=dw32, a_mn=True, b_mn=True), lambda i: torch.mm(dys[i].t(), xs[i], out=dw16))):
        t_o, t_r = timed(ours), timed(ref)
        row[op] = {"ours_us": round(t_o, 1), "torch_us": round(t_r, 1), "ours_tflops": round(fl / t_o / 1e6, 0), "torch_tflops": round(fl / t_r / 1e6, 0)}
    out[name] = row
    print(name, json.dumps(row), flush=True)
print(json.dumps(out))
