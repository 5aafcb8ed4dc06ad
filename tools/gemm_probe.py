"""Time the tcgen05 GEMM on the MLP's shapes against torch (cuBLAS) fp32 and tf32.

Each case is captured in a CUDA graph of ROT launches that rotate over ROT distinct operand sets
(total footprint > the 126 MB L2, so every launch streams its big operand from HBM), timed with
CUDA events around graph replays after warm-up. Usage: python tools/gemm_probe.py [batch]"""
import json
import sys

import torch

from ray_torch_distributed_checkpoint_b200 import ops

B = int(sys.argv[1]) if len(sys.argv) > 1 else 16384
dev = "cuda"
ROT = 6


def graph_time(fns, reps=5):
    """fns: list of ROT callables (one per operand set). Returns ms per launch."""
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for f in fns:
            f()
        s.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g, stream=s):
            for f in fns:
                f()
        for _ in range(3):
            g.replay()
        s.synchronize()
        ts = []
        for _ in range(reps):
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            a.record(s); g.replay(); b.record(s); s.synchronize()
            ts.append(a.elapsed_time(b) / len(fns))
    ts.sort()
    return ts[len(ts) // 2]


xs = [torch.randn(B, 784, device=dev) for _ in range(ROT)]
hs = [torch.relu(torch.randn(B, 512, device=dev)) for _ in range(ROT)]
dys = [torch.randn(B, 512, device=dev) for _ in range(ROT)]
outs = [torch.empty(B, 512, device=dev) for _ in range(ROT)]
w1 = torch.randn(512, 784, device=dev) * .03; b1 = torch.randn(512, device=dev)
w2 = torch.randn(512, 512, device=dev) * .03
dw1 = torch.empty(512, 784, device=dev); dw2 = torch.empty(512, 512, device=dev)
res = {"batch": B, "rot": ROT}
cases = {
    "fwd1": (2 * B * 784 * 512,
             lambda i, prec: ops.linear_forward(xs[i], w1, b1, relu=True, drop_p=0.25, seed=1, out=outs[i], precision=prec),
             lambda i: torch.dropout(torch.relu(torch.addmm(b1, xs[i], w1.T)), 0.25, True)),
    "fwd1_nodrop": (2 * B * 784 * 512,
             lambda i, prec: ops.linear_forward(xs[i], w1, b1, relu=True, out=outs[i], precision=prec),
             lambda i: torch.relu(torch.addmm(b1, xs[i], w1.T))),
    "fwd2": (2 * B * 512 * 512,
             lambda i, prec: ops.linear_forward(hs[i], w2, b1, relu=True, drop_p=0.25, seed=1, out=outs[i], precision=prec),
             lambda i: torch.dropout(torch.relu(torch.addmm(b1, hs[i], w2.T)), 0.25, True)),
    "dgrad2": (2 * B * 512 * 512,
               lambda i, prec: ops.linear_dgrad(dys[i], w2, mask_src=hs[i], mask_scale=4 / 3, out=outs[i], precision=prec),
               lambda i: (dys[i] @ w2) * (hs[i] > 0) * (4 / 3)),
    "wgrad1": (2 * B * 784 * 512, lambda i, prec: ops.linear_wgrad(dys[i], xs[i], out=dw1, precision=prec), lambda i: dys[i].T @ xs[i]),
    "wgrad2": (2 * B * 512 * 512, lambda i, prec: ops.linear_wgrad(dys[i], hs[i], out=dw2, precision=prec), lambda i: dys[i].T @ hs[i]),
}
for name, (flops, mine, ref) in cases.items():
    r = {}
    for prec in (3, 1):
        t = graph_time([(lambda i=i: mine(i, prec)) for i in range(ROT)])
        r[f"rtdc_p{prec}_ms"] = round(t, 4); r[f"rtdc_p{prec}_tflops"] = round(flops / t / 1e9, 1)
    torch.backends.cuda.matmul.allow_tf32 = False
    t = graph_time([(lambda i=i: ref(i)) for i in range(ROT)]); r["torch_fp32_ms"] = round(t, 4); r["torch_fp32_tflops"] = round(flops / t / 1e9, 1)
    torch.backends.cuda.matmul.allow_tf32 = True
    t = graph_time([(lambda i=i: ref(i)) for i in range(ROT)]); r["torch_tf32_ms"] = round(t, 4)
    torch.backends.cuda.matmul.allow_tf32 = False
    res[name] = r
    print(name, json.dumps(r), flush=True)
json.dump(res, open("gpurun_out/gemm_probe.json", "w"), indent=1)
