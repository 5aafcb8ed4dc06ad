"""bf16 LayerNorm forward / backward and bf16 column sums (csrc/layernorm.cu) against plain PyTorch fp32 references."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("rows,C,bias", [(16384, 768, True), (1000, 768, False), (37, 64, True), (513, 1024, True), (5, 8, True), (300, 520, True)])
def test_layernorm_bf16_matches_fp32_reference(rows, C, bias):
    from ray_torch_distributed_checkpoint_b200.ops import layer_norm_bf16
    torch.manual_seed(rows + C)
    x = (torch.randn(rows, C, device="cuda") * 2.0 + 0.5).to(torch.bfloat16).requires_grad_(True)
    w = (1.0 + 0.2 * torch.randn(C, device="cuda")).to(torch.bfloat16).requires_grad_(True)
    b = (0.1 * torch.randn(C, device="cuda")).to(torch.bfloat16).requires_grad_(True) if bias else None
    dy = torch.randn(rows, C, device="cuda").to(torch.bfloat16)
    y = layer_norm_bf16(x, w, b, 1e-5)
    assert y.dtype == torch.bfloat16 and y.shape == x.shape
    y.backward(dy)

    x32, w32 = x.detach().float().requires_grad_(True), w.detach().float().requires_grad_(True)
    b32 = b.detach().float().requires_grad_(True) if bias else None
    ref = F.layer_norm(x32, (C,), w32, b32, 1e-5)
    ref.backward(dy.float())
    assert (y.float() - ref).abs().max().item() < 2e-2 * ref.abs().max().item()
    assert (x.grad.float() - x32.grad).abs().max().item() < 2e-2 * x32.grad.abs().max().item() + 1e-3
    # column sums over `rows` rows in fp32, rounded once to bf16
    assert (w.grad.float() - w32.grad).abs().max().item() < 6e-3 * w32.grad.abs().max().item() + 1e-2
    if bias:
        assert (b.grad.float() - b32.grad).abs().max().item() < 6e-3 * b32.grad.abs().max().item() + 1e-2


def test_layernorm_module_swap_keeps_state_dict_and_3d_inputs():
    from ray_torch_distributed_checkpoint_b200.ops import LayerNormBf16, use_fused_layernorm
    torch.manual_seed(0)
    ref = torch.nn.Sequential(torch.nn.LayerNorm(256), torch.nn.Linear(256, 256), torch.nn.LayerNorm(256)).cuda().to(torch.bfloat16)
    ours = torch.nn.Sequential(torch.nn.LayerNorm(256), torch.nn.Linear(256, 256), torch.nn.LayerNorm(256)).cuda().to(torch.bfloat16)
    ours.load_state_dict(ref.state_dict())
    use_fused_layernorm(ours)
    assert isinstance(ours[0], LayerNormBf16) and list(ours.state_dict()) == list(ref.state_dict())
    x = torch.randn(3, 50, 256, device="cuda", dtype=torch.bfloat16)
    xa, xb = x.clone().requires_grad_(True), x.clone().requires_grad_(True)
    ya, yb = ref(xa), ours(xb)
    ya.float().square().mean().backward()
    yb.float().square().mean().backward()
    assert (ya.float() - yb.float()).abs().max().item() < 3e-2
    assert (xa.grad.float() - xb.grad.float()).abs().max().item() < 3e-2 * xa.grad.float().abs().max().item() + 1e-5
    for pa, pb in zip(ref.parameters(), ours.parameters()):
        assert (pa.grad.float() - pb.grad.float()).abs().max().item() < 3e-2 * pa.grad.float().abs().max().item() + 1e-4


@pytest.mark.parametrize("rows,cols", [(16384, 768), (16384, 2304), (100, 8), (777, 3072), (63, 264)])
def test_colsum_bf16_matches_fp32_sum(rows, cols):
    from ray_torch_distributed_checkpoint_b200.ops import colsum_bf16
    torch.manual_seed(rows + cols)
    canvas = torch.randn(rows, cols + 8, device="cuda").to(torch.bfloat16)
    x = canvas[:, :cols]                                       # row pitch larger than the width
    ref = x.float().sum(0)
    out32 = colsum_bf16(x, out_dtype=torch.float32)
    assert (out32 - ref).abs().max().item() < 1e-4 * (x.float().abs().sum(0).max().item() + 1.0)
    out16 = colsum_bf16(x)
    assert out16.dtype == torch.bfloat16
    assert (out16.float() - ref).abs().max().item() < 5e-3 * ref.abs().max().item() + 1e-2


def test_linear_bf16_bias_gradient_uses_colsum():
    from ray_torch_distributed_checkpoint_b200.ops import LinearBf16
    torch.manual_seed(2)
    lin = torch.nn.Linear(128, 256).cuda().to(torch.bfloat16)
    ours = LinearBf16(lin)
    x = torch.randn(512, 128, device="cuda", dtype=torch.bfloat16)
    dy = torch.randn(512, 256, device="cuda", dtype=torch.bfloat16)
    ours(x).backward(dy)
    assert (lin.bias.grad.float() - dy.float().sum(0)).abs().max().item() < 5e-3 * dy.float().sum(0).abs().max().item() + 1e-2
