"""Run one kernel case a few times (for ncu / sanitizer captures). Usage:
python tools/profile_case.py <fwd1|fwd2|dgrad2|wgrad1|wgrad2> <precision> [batch] [iters]"""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))

import sys

import torch

from ray_torch_distributed_checkpoint_b200 import ops

case, prec = sys.argv[1], int(sys.argv[2])
B = int(sys.argv[3]) if len(sys.argv) > 3 else 16384
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 6
dev = "cuda"
x = torch.randn(B, 784, device=dev); h = torch.relu(torch.randn(B, 512, device=dev)); dy = torch.randn(B, 512, device=dev)
w1 = torch.randn(512, 784, device=dev) * .03; w2 = torch.randn(512, 512, device=dev) * .03; b1 = torch.randn(512, device=dev)
out = torch.empty(B, 512, device=dev); dw1 = torch.empty(512, 784, device=dev); dw2 = torch.empty(512, 512, device=dev)
fns = {
    "fwd1": lambda: ops.linear_forward(x, w1, b1, relu=True, drop_p=0.25, seed=1, out=out, precision=prec),
    "fwd1_nodrop": lambda: ops.linear_forward(x, w1, b1, relu=True, out=out, precision=prec),
    "fwd2": lambda: ops.linear_forward(h, w2, b1, relu=True, drop_p=0.25, seed=1, out=out, precision=prec),
    "dgrad2": lambda: ops.linear_dgrad(dy, w2, mask_src=h, mask_scale=4 / 3, out=out, precision=prec),
    "wgrad1": lambda: ops.linear_wgrad(dy, x, out=dw1, precision=prec),
    "wgrad2": lambda: ops.linear_wgrad(dy, h, out=dw2, precision=prec),
}
for _ in range(iters):
    fns[case]()
torch.cuda.synchronize()
print("ok", case, prec, B)
