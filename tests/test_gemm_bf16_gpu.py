"""bf16 tcgen05 GEMM against a plain PyTorch fp32 reference of the same op (bf16 inputs widened to fp32), all four operand
major combinations, tails in every dimension, epilogues, split-K, and the Linear layer's three GEMMs under autograd."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(a, b, a_mn, b_mn):
    A = a.float().t() if a_mn else a.float()
    B = b.float().t() if b_mn else b.float()
    return A @ B.t()


@pytest.mark.parametrize("a_mn,b_mn", [(False, False), (False, True), (True, False), (True, True)])
@pytest.mark.parametrize("M,N,K", [(256, 256, 256), (384, 768, 1024), (200, 136, 72), (1000, 2304, 768), (128, 64, 64)])
def test_gemm_bf16_matches_fp32_reference(a_mn, b_mn, M, N, K):
    from ray_torch_distributed_checkpoint_b200.ops import gemm_bf16
    torch.manual_seed(M + N + K)
    a = (torch.randn(K, M) if a_mn else torch.randn(M, K)).cuda().to(torch.bfloat16)
    b = (torch.randn(K, N) if b_mn else torch.randn(N, K)).cuda().to(torch.bfloat16)
    ref = _ref(a, b, a_mn, b_mn)
    scale = ref.abs().max().item()
    out32 = gemm_bf16(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32, splits=1)
    assert (out32 - ref).abs().max().item() < 2e-4 * scale + 1e-3, (out32 - ref).abs().max().item()
    out16 = gemm_bf16(a, b, a_mn=a_mn, b_mn=b_mn)
    assert out16.dtype == torch.bfloat16
    assert (out16.float() - ref).abs().max().item() < 8e-3 * scale + 1e-2
    if K >= 256:
        split = gemm_bf16(a, b, a_mn=a_mn, b_mn=b_mn, out_dtype=torch.float32, splits=3)
        assert (split - ref).abs().max().item() < 2e-4 * scale + 1e-3


def test_gemm_bf16_epilogues_and_padded_output():
    from ray_torch_distributed_checkpoint_b200.ops import gemm_bf16
    torch.manual_seed(0)
    M, N, K = 300, 520, 192
    a = torch.randn(M, K).cuda().to(torch.bfloat16)
    b = (torch.randn(N, K) * 0.1).cuda().to(torch.bfloat16)
    bias = torch.randn(N).cuda().to(torch.bfloat16)
    pre = 0.5 * (a.float() @ b.float().t()) + bias.float()
    for act, fn in ((0, lambda t: t), (1, torch.relu), (2, lambda t: torch.nn.functional.gelu(t, approximate="tanh")),
                    (3, torch.nn.functional.gelu)):
        canvas = torch.full((M + 3, N + 8), 7.0, device="cuda", dtype=torch.bfloat16)
        out = gemm_bf16(a, b, out=canvas, bias=bias, act=act, alpha=0.5)
        assert (out[:M, :N].float() - fn(pre)).abs().max().item() < 3e-2
        assert (canvas[M:] == 7).all() and (canvas[:, N:] == 7).all()          # nothing written outside [M, N]
    with pytest.raises(Exception):
        gemm_bf16(a, b, bias=bias, out_dtype=torch.float32, splits=4)              # split-K cannot carry a bias


def test_linear_bf16_under_autograd_matches_torch():
    from ray_torch_distributed_checkpoint_b200.ops import LinearBf16, use_bf16_linears
    torch.manual_seed(1)
    ref = torch.nn.Sequential(torch.nn.Linear(768, 3072), torch.nn.GELU(approximate="tanh"), torch.nn.Linear(3072, 768)).cuda().to(torch.bfloat16)
    ours = torch.nn.Sequential(torch.nn.Linear(768, 3072), torch.nn.GELU(approximate="tanh"), torch.nn.Linear(3072, 768)).cuda().to(torch.bfloat16)
    ours.load_state_dict(ref.state_dict())
    use_bf16_linears(ours)
    assert isinstance(ours[0], LinearBf16) and list(ours.state_dict()) == list(ref.state_dict())
    x = torch.randn(4, 512, 768, device="cuda", dtype=torch.bfloat16)
    xr, xo = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    yr, yo = ref(xr), ours(xo)
    assert (yr.float() - yo.float()).abs().max().item() < 3e-2 * yr.float().abs().max().item() + 1e-2
    g = torch.randn_like(yr)
    yr.backward(g); yo.backward(g)
    rel = lambda u, v: ((u.float() - v.float()).norm() / v.float().norm()).item()
    assert rel(xo.grad, xr.grad) < 2e-2
    for (n, po), (_, pr) in zip(ours.named_parameters(), ref.named_parameters()):
        assert rel(po.grad, pr.grad) < 2e-2, n


def test_adamw_bf16_arena_matches_fp32_adamw_on_master_weights():
    from ray_torch_distributed_checkpoint_b200.optim import AdamW
    torch.manual_seed(2)
    net = torch.nn.Sequential(torch.nn.Linear(64, 96), torch.nn.Linear(96, 33)).cuda().to(torch.bfloat16)
    ref = [p.detach().float().clone().requires_grad_(True) for p in net.parameters()]
    opt_r = torch.optim.AdamW(ref, lr=1e-2, weight_decay=0.01)
    opt_o = AdamW(net, lr=1e-2)
    assert opt_o.bf16 and opt_o.master.dtype == torch.float32
    for step in range(5):
        grads = [torch.randn_like(p).to(torch.bfloat16) for p in net.parameters()]
        opt_o.zero_grad()
        for p, g, r in zip(net.parameters(), grads, ref):
            p.grad.copy_(g)
            r.grad = g.float()
        opt_o.step(); opt_r.step()
    for p, r in zip(net.parameters(), ref):
        assert torch.allclose(p.float(), r.to(torch.bfloat16).float(), rtol=1e-2, atol=1e-3)    # bf16 copy of the same fp32 trajectory
    for i, r in enumerate(ref):                                                               # the master copy itself is fp32-exact
        off, n, _ = opt_o._slot[i]
        assert torch.allclose(opt_o.master[off:off + n], r.detach().reshape(-1), rtol=1e-5, atol=1e-6)
