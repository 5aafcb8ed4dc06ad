"""LM-head loss entry point off the GPU: the plain-PyTorch route gives F.cross_entropy's value and gradients, and a GPT-2
with ``fused_head`` set behaves like the stock one."""
import torch
import torch.nn.functional as F


def test_lm_head_cross_entropy_cpu_fallback_matches_torch():
    from ray_torch_distributed_checkpoint_b200.ops.lm_head import lm_head_cross_entropy
    torch.manual_seed(0)
    h = torch.randn(3, 5, 16, requires_grad=True)
    w = torch.randn(33, 16, requires_grad=True)
    tgt = torch.randint(0, 33, (3, 5))
    tgt[1, 2] = -100
    loss = lm_head_cross_entropy(h, w, tgt)
    ref = F.cross_entropy((h.detach().reshape(-1, 16) @ w.detach().t()), tgt.reshape(-1))
    assert torch.allclose(loss.detach(), ref, atol=1e-6)
    loss.backward()
    assert h.grad is not None and w.grad is not None and h.grad.shape == h.shape


def test_gpt2_fused_head_flag_is_harmless_on_cpu():
    from ray_torch_distributed_checkpoint_b200.models import GPT2, GPT2Config
    torch.manual_seed(1)
    cfg = GPT2Config(vocab_size=101, n_positions=16, n_embd=32, n_layer=1, n_head=2)
    m = GPT2(cfg)
    idx = torch.randint(0, 101, (2, 16))
    tgt = torch.randint(0, 101, (2, 16))
    plain = m(idx, tgt)
    m.fused_head = True
    fused = m(idx, tgt)
    assert torch.allclose(plain, fused, atol=1e-6)
    assert m(idx).shape == (2, 16, 101)              # logits path untouched


def test_layer_norm_entry_point_falls_back_on_cpu():
    from ray_torch_distributed_checkpoint_b200.ops.layernorm import layer_norm_bf16, use_fused_layernorm
    torch.manual_seed(2)
    ln = torch.nn.LayerNorm(24)
    x = torch.randn(4, 6, 24)
    assert torch.allclose(layer_norm_bf16(x, ln.weight, ln.bias, ln.eps), ln(x), atol=1e-6)
    seq = torch.nn.Sequential(torch.nn.LayerNorm(24), torch.nn.Linear(24, 24))
    keys = list(seq.state_dict())
    want = seq(x)
    use_fused_layernorm(seq)
    assert list(seq.state_dict()) == keys and torch.allclose(seq(x), want, atol=1e-6)
