"""One launch each of the kernels this framework puts on the GPT-2 small step (for ncu): the bf16 tcgen05 CTA-pair GEMM on
the c_attn shapes (forward / dgrad / wgrad) and on the LM head (forward, odd vocabulary), the fused cross-entropy over the
head's logits, LayerNorm forward / backward (+ the partial reductions), the bias-gradient column sums, bf16 AdamW.
Two iterations; capture the second (`-s 14 -c 14`)."""
import os as _os, sys as _sys
_sys.path.insert(0, _os.path.dirname(_os.path.dirname(_os.path.abspath(__file__))))

import torch

from ray_torch_distributed_checkpoint_b200 import ops

dev = "cuda"
torch.manual_seed(0)
M, C, V = 16384, 768, 50257
x = torch.randn(M, C, device=dev).to(torch.bfloat16)
w = (torch.randn(3 * C, C, device=dev) * 0.02).to(torch.bfloat16)
b = torch.zeros(3 * C, device=dev, dtype=torch.bfloat16)
dy = torch.randn(M, 3 * C, device=dev).to(torch.bfloat16)
wte = (torch.randn(V, C, device=dev) * 0.02).to(torch.bfloat16)
ld = (V + 7) // 8 * 8
logits_buf = torch.empty(M, ld, device=dev, dtype=torch.bfloat16)
logits = logits_buf[:, :V]
tgt = torch.randint(0, V, (M,), device=dev)
scale = torch.full((1,), 1.0 / M, device=dev)
gamma = torch.ones(C, device=dev, dtype=torch.bfloat16)
beta = torch.zeros(C, device=dev, dtype=torch.bfloat16)
n = 4 << 20
p16 = torch.randn(n, device=dev).to(torch.bfloat16); g16 = torch.randn(n, device=dev).to(torch.bfloat16)
master, m_, v_ = p16.float(), torch.zeros(n, device=dev), torch.zeros(n, device=dev)
adam = ops.AdamState(dev, 3e-4, (0.9, 0.999), 1e-8, 0.01, True)
for _ in range(2):
    y = ops.gemm_bf16(x, w, bias=b)                                              # 1: forward  16384 x 768 -> 2304
    dx = ops.gemm_bf16(dy, w, b_mn=True)                                         # 2: dgrad    16384 x 2304 -> 768
    dw = ops.gemm_bf16(dy, x, a_mn=True, b_mn=True, out_dtype=torch.float32)     # 3: wgrad    2304 x 768, K = 16384 (split)
    ops.gemm_bf16(x, wte, out=logits)                                            # 4: LM head  16384 x 768 -> 50257
    rl = ops.softmax_ce_bf16_(logits, tgt, scale)                                # 5: cross-entropy + gradient in place
    xin = x.clone().requires_grad_(True)
    yn = ops.layer_norm_bf16(xin, gamma.requires_grad_(True), beta.requires_grad_(True))   # 6: LayerNorm forward
    yn.backward(x)                                                               # 7-9: LayerNorm backward + 2 partial reductions
    db = ops.colsum_bf16(dy)                                                     # 10-11: bias gradient + partial reduction
    ops.adam_step(p16, g16, m_, v_, adam, master=master)                         # 12-13: step counter + AdamW
torch.cuda.synchronize()
print("ok", float(rl.mean()), float(db.float().abs().mean()))
