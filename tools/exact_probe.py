"""Debug probe: isolates the exact-operand / alpha / row_bias features of ops.gemm."""
import itertools
import torch
from ray_torch_distributed_checkpoint_b200 import ops

torch.manual_seed(3)
ints = lambda *shape: torch.randint(0, 256, shape, device="cuda").float()
for (M, N, K) in ((300, 520, 784), (512, 512, 784), (128, 256, 64)):
    for a_ex, b_ex, use_alpha, use_rb, use_bias, splits in itertools.product((0, 1), (0, 1), (0, 1), (0, 1), (0, 1), (1, 0)):
        a = ints(M, K) if a_ex else torch.randn(M, K, device="cuda")
        b = ints(N, K) if b_ex else torch.randn(N, K, device="cuda")
        bias = torch.randn(N, device="cuda") if use_bias else None
        rb = torch.randn(M, device="cuda") if use_rb else None
        alpha = 2.0 / 255 if use_alpha else 1.0
        out = torch.empty(M, N, device="cuda")
        ops.gemm(a, b, out, bias=bias, a_exact=bool(a_ex), b_exact=bool(b_ex), alpha=alpha, row_bias=rb, row_bias_scale=-1.0, splits=splits)
        ref = alpha * (a.double() @ b.double().T)
        if use_bias: ref = ref + bias.double()
        if use_rb: ref = ref - rb.double()[:, None]
        scale = (alpha * (a.double().abs() @ b.double().abs().T)).max().item()
        err = (out.double() - ref).abs().max().item() / scale
        flag = "" if err < 1.2e-5 else "   <-- BAD"
        if flag or (a_ex + b_ex + use_alpha + use_rb == 0):
            print(f"M{M} N{N} K{K} a_ex={a_ex} b_ex={b_ex} alpha={use_alpha} rb={use_rb} bias={use_bias} splits={splits} plan={ops.gemm_plan(M, N, K, splits=splits)['splits']} relerr={err:.3e}{flag}")
print("done")
