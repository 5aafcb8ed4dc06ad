"""A few launches each of the kernels that are not part of the MLP training step (for ncu): the bf16 tcgen05 GEMM on GPT-2
shapes (forward / dgrad / wgrad), the resident loader's gather kernel, the evaluation loss kernel, the activation backward,
the bf16 AdamW kernel, and the single-rank instances of the collective kernels (two-shot, fused optimizer step: same code,
no peers)."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))

import torch

from ray_torch_distributed_checkpoint_b200 import ops
from ray_torch_distributed_checkpoint_b200.parallel.symm import SymmetricArena

dev = "cuda"
torch.manual_seed(0)
M, C = 16384, 768
x = torch.randn(M, C, device=dev).to(torch.bfloat16)
w = (torch.randn(3 * C, C, device=dev) * 0.02).to(torch.bfloat16)
b = torch.zeros(3 * C, device=dev, dtype=torch.bfloat16)
dy = torch.randn(M, 3 * C, device=dev).to(torch.bfloat16)
imgs = torch.randint(0, 256, (60000, 784), dtype=torch.uint8, device=dev)
labels = torch.randint(0, 10, (60000,), dtype=torch.uint8, device=dev)
idx = torch.randperm(60000, device=dev)[:18944].to(torch.int32)
xb = torch.empty(18944, 784, dtype=torch.uint8, device=dev); yb = torch.empty(18944, dtype=torch.int64, device=dev)
z = torch.randn(18944, 16, device=dev); yl = torch.randint(0, 10, (18944,), device=dev)
g = torch.randn(18944, 512, device=dev); h = torch.relu(torch.randn(18944, 512, device=dev))
n = 4 << 20
p16 = torch.randn(n, device=dev).to(torch.bfloat16); g16 = torch.randn(n, device=dev).to(torch.bfloat16)
master, m_, v_ = p16.float(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
adam = ops.AdamState(dev, 3e-4, (0.9, 0.999), 1e-8, 0.01, True)
arena = SymmetricArena(16 << 20, device=torch.device("cuda", 0))
N_ARENA = 669728
g_off, ga = arena.alloc(N_ARENA); p_off, pa = arena.alloc(N_ARENA); m_off, ma = arena.alloc(N_ARENA)
ga.fill_(1e-3); pa.fill_(0.5); ma.zero_()
sgd = ops.SgdState(dev, 1e-3, 0.9)
counter = torch.zeros(1, dtype=torch.int32, device=dev)
for _ in range(3):
    y = ops.gemm_bf16(x, w, bias=b)                                              # forward  16384 x 768 -> 2304
    dx = ops.gemm_bf16(dy, w, b_mn=True)                                         # dgrad    16384 x 2304 -> 768
    dw = ops.gemm_bf16(dy, x, a_mn=True, b_mn=True, out_dtype=torch.float32)     # wgrad    2304 x 768, K = 16384 (split)
    ops.u8_gather_rows(imgs, idx, xb, labels=labels, y_out=yb)
    ops.softmax_cross_entropy(z, yl, classes=10)
    ops.act_backward(g, h, scale=4 / 3)
    ops.adam_step(p16, g16, m_, v_, adam, master=master)
    arena.all_reduce(g_off, N_ARENA, scale=1.0, algo="two_shot")
    arena.sgd_step(g_off, p_off, m_off, N_ARENA, sgd, counter=counter, algo="two_shot")
torch.cuda.synchronize()
ref = x.float() @ w.float().t()
print("ok", float((y.float() - ref).abs().max()), float(dw.abs().mean()))
arena.close()
