"""Per-CTA timeline of the CTA-pair GEMM kernel from its RTDC_GEMM_DEBUG_BUF stamps ([grid][16] clock64 values)."""
# NOTE: needs a library built with the debug hooks: RTDC_EXTRA_NVCC_FLAGS=-DRTDC_GEMM_DEBUG_HOOKS python -m ray_torch_distributed_checkpoint_b200.ops._build --force

import os, sys
import torch
from ray_torch_distributed_checkpoint_b200 import ops
B = 18944; dev = "cuda"
x = torch.randint(0, 256, (B, 784), device=dev).float(); w1 = torch.randn(512, 784, device=dev) * .03; b1 = torch.randn(512, device=dev)
h = torch.relu(torch.randn(B, 512, device=dev)); w2 = torch.randn(512, 512, device=dev) * .03
dy = torch.randn(B, 512, device=dev); out = torch.empty(B, 512, device=dev)
buf = torch.zeros(512 * 16, dtype=torch.int64, device=dev)
part = torch.zeros(4 * 148, 512, device=dev)
cases = {"fwd1": lambda: ops.gemm(x, w1, out, bias=b1, relu=True, drop_p=0.25, seed=1),
         "fwd1_exact": lambda: ops.gemm(x, w1, out, bias=b1, relu=True, drop_p=0.25, seed=1, a_exact=True),
         "fwd1_exact_nodrop": lambda: ops.gemm(x, w1, out, bias=b1, relu=True, a_exact=True),
         "fwd2": lambda: ops.gemm(h, w2, out, bias=b1, relu=True, drop_p=0.25, seed=1),
         "fwd2_bn128": lambda: ops.gemm(h, w2, out, bias=b1, relu=True, drop_p=0.25, seed=1, bn=128),
         "fwd2_bn64": lambda: ops.gemm(h, w2, out, bias=b1, relu=True, drop_p=0.25, seed=1, bn=64),
         "dgrad2": lambda: ops.linear_dgrad(dy, w2, mask_src=h, mask_scale=4 / 3, out=out),
         "dgrad2_colsum": lambda: ops.linear_dgrad(dy, w2, mask_src=h, mask_scale=4 / 3, out=out, colsum_partial=part),
         "dgrad2_nomask": lambda: ops.linear_dgrad(dy, w2, out=out)}
for case in (sys.argv[1:] or list(cases)):
    os.environ.pop("RTDC_GEMM_DEBUG_BUF", None)
    for _ in range(3): cases[case]()
    torch.cuda.synchronize(); buf.zero_()
    os.environ["RTDC_GEMM_DEBUG_BUF"] = str(buf.data_ptr())
    cases[case](); torch.cuda.synchronize()
    os.environ.pop("RTDC_GEMM_DEBUG_BUF", None)
    t = buf.view(-1, 16).cpu(); n = int((t[:, 7] != 0).sum()); t = t[:n].double()
    c0 = t[:, 0:1]
    rel = (t[:, :10] - c0) / 1965.0          # us at the nominal 1965 MHz
    lead = t[:, 8] != 0
    f = lambda col, m=None: f"{rel[:, col][m].mean():.1f}" if m is not None else f"{rel[:, col].mean():.1f}"
    print(f"{case}: {n} CTAs span {(t[:, 6].max() - t[:, 7].min()) / 1000:.1f} us | epi0 {f(1)}->{f(2)}  epiLast {f(3)}->{f(4)}  end {f(5)} | "
          f"issuer first unit done {f(8, lead)} last {f(9, lead)}", flush=True)
    n_it = t[lead][:, 13].mean()
    print(f"    issuer, last unit, cycles per stage over {n_it:.0f} stages: wait ready {t[lead][:, 10].mean() / n_it:.0f}  "
          f"issue MMAs {t[lead][:, 11].mean() / n_it:.0f}  commit {t[lead][:, 12].mean() / n_it:.0f}", flush=True)
