"""Per-CTA phase timeline of the GEMM kernel (RTDC_GEMM_DEBUG_BUF stamps). Usage: gemm_stamps.py [case ...]"""
import os, sys
import torch
from ray_torch_distributed_checkpoint_b200 import ops
B = 16384; dev = "cuda"
x = torch.randn(B, 784, device=dev); w1 = torch.randn(512, 784, device=dev) * .03; b1 = torch.randn(512, device=dev)
h = torch.relu(torch.randn(B, 512, device=dev)); dy = torch.randn(B, 512, device=dev)
out = torch.empty(B, 512, device=dev); dw1 = torch.empty(512, 784, device=dev)
buf = torch.zeros(4096 * 8, dtype=torch.int64, device=dev)
cases = {
    "fwd1": lambda prec: ops.gemm(x, w1, out, bias=b1, relu=True, precision=prec, bn=256),
    "wgrad1": lambda prec: ops.linear_wgrad(dy, x, out=dw1, precision=prec),
    "dgrad2": lambda prec: ops.linear_dgrad(dy, w1[:, :512].contiguous(), mask_src=h, out=out, precision=prec),
}
for case in (sys.argv[1:] or ["fwd1"]):
    for prec in (1, 3):
        for dbg in (0, 1, 2):
            os.environ["RTDC_GEMM_DEBUG"] = str(dbg)
            os.environ.pop("RTDC_GEMM_DEBUG_BUF", None)
            for _ in range(3): cases[case](prec)
            torch.cuda.synchronize()
            buf.zero_()
            os.environ["RTDC_GEMM_DEBUG_BUF"] = str(buf.data_ptr())
            cases[case](prec)
            torch.cuda.synchronize()
            t = buf.view(-1, 8).cpu()
            n = int((t[:, 7] != 0).sum())
            t = t[:n]
            g0 = t[:, 7].min()
            print(f"{case} prec={prec} debug={dbg}: {n} CTAs, kernel span {(t[:,6].max()-g0).item()/1000:.1f} us")
            if dbg == 0:
                for b in sorted(set([0, n // 2, n - 1])):
                    r = t[b]
                    print(f"  cta {b:3d}: start@{(r[7]-g0).item()/1000:6.1f}us epi_wait_start {r[1]-r[0]:6d} mma_issued {r[2]-r[0]:6d} accum_ready {r[3]-r[0]:6d} epi_done {r[4]-r[0]:6d} cyc  life {(r[6]-r[7]).item()/1000:.1f} us")
