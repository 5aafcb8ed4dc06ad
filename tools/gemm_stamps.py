"""Per-CTA lifetime of the (persistent) GEMM kernel from its RTDC_GEMM_DEBUG_BUF stamps."""
# NOTE: needs a library built with the debug hooks: RTDC_EXTRA_NVCC_FLAGS=-DRTDC_GEMM_DEBUG_HOOKS python -m ray_torch_distributed_checkpoint_b200.ops._build --force

import os, sys
import torch
from ray_torch_distributed_checkpoint_b200 import ops
B = 18944; dev = "cuda"
x = torch.randn(B, 784, device=dev); w1 = torch.randn(512, 784, device=dev) * .03; b1 = torch.randn(512, device=dev)
dy = torch.randn(B, 512, device=dev); out = torch.empty(B, 512, device=dev); dw1 = torch.empty(512, 784, device=dev)
buf = torch.zeros(4096 * 8, dtype=torch.int64, device=dev)
cases = {"fwd1": lambda prec: ops.gemm(x, w1, out, bias=b1, relu=True, precision=prec),
         "wgrad1": lambda prec: ops.linear_wgrad(dy, x, out=dw1, precision=prec)}
for case in (sys.argv[1:] or ["fwd1"]):
    for prec in (1, 3):
        os.environ.pop("RTDC_GEMM_DEBUG_BUF", None)
        for _ in range(3): cases[case](prec)
        torch.cuda.synchronize(); buf.zero_()
        os.environ["RTDC_GEMM_DEBUG_BUF"] = str(buf.data_ptr())
        cases[case](prec); torch.cuda.synchronize()
        t = buf.view(-1, 8).cpu(); n = int((t[:, 7] != 0).sum()); t = t[:n]; g0 = t[:, 7].min()
        life = (t[:, 6] - t[:, 7]).float() / 1000
        print(f"{case} prec={prec}: {n} CTAs, span {(t[:,6].max()-g0).item()/1000:.1f} us, CTA life mean {life.mean():.1f} max {life.max():.1f} us, first unit MMAs issued after {(t[:,2]-t[:,0]).float().mean():.0f} cyc")
