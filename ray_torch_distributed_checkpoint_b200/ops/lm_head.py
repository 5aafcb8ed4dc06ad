"""Fused LM head: ``cross_entropy(h @ W^T, targets)`` for a wide vocabulary without any full-size fp32 tensor.

Serves the GPT-2 small configuration of BASELINE.json (config #5); not a reference capability (the reference's loss is the
10-class ``nn.CrossEntropyLoss`` of my_ray_module.py:141). Stock PyTorch's ``F.cross_entropy(logits.float(), targets)`` moves
~30 GB through HBM per step for 16 x 1024 tokens x 50 257 classes (upcast, log-softmax, its saved output, and the gradient
of each); here the bf16 logits come out of the tcgen05 GEMM (``csrc/gemm_bf16.cu``), ONE kernel pass
(``csrc/lm_head.cu``) turns them in place into the loss gradient, and the head's dgrad / wgrad GEMMs read that buffer as stored.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn.functional as F

from . import _lib
from .gemm_bf16 import gemm_bf16


def _lib_lm():
    lib = _lib.load()
    if not getattr(lib, "_lm_declared", False):
        lib.rtdc_lm_head_ce_bf16.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_int, C.c_int, C.c_longlong, C.c_void_p,
                                             C.c_void_p, C.c_void_p]
        lib.rtdc_lm_head_ce_bf16.restype = C.c_int
        lib._lm_declared = True
    return lib


def softmax_ce_bf16_(logits: torch.Tensor, targets: torch.Tensor, grad_scale: torch.Tensor, ignore_index: int = -100,
                     stream=None) -> torch.Tensor:
    """Overwrite the bf16 ``logits`` ``[rows, V]`` (unit inner stride, 16-byte aligned rows) with
    ``(softmax - onehot) * grad_scale`` and return the per-row losses ``[rows]`` (fp32, natural log; 0 where the target is
    ``ignore_index``). ``grad_scale`` is a device scalar (fp32)."""
    if not logits.is_cuda or logits.dtype != torch.bfloat16 or logits.dim() != 2 or logits.stride(1) != 1:
        raise ValueError("logits must be a 2-D bfloat16 CUDA tensor with unit inner stride")
    if targets.dtype != torch.int64 or targets.numel() != logits.shape[0] or not targets.is_contiguous():
        raise ValueError("targets must be a contiguous int64 vector with one entry per row")
    if grad_scale.dtype != torch.float32 or grad_scale.numel() != 1 or not grad_scale.is_cuda:
        raise ValueError("grad_scale must be a CUDA float32 scalar")
    rows, vocab = logits.shape
    row_loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
    _lib.check(_lib_lm().rtdc_lm_head_ce_bf16(logits.data_ptr(), logits.stride(0), targets.data_ptr(), rows, vocab, int(ignore_index),
                                              grad_scale.data_ptr(), row_loss.data_ptr(), _lib.stream_ptr(stream)))
    return row_loss


class _LMHeadCEFn(torch.autograd.Function):
    """loss = mean over counted rows of CE(h W^T, targets). The gradient w.r.t. the logits is formed in the forward pass (in the
    logits' own buffer), so backward is two GEMMs."""

    @staticmethod
    def forward(ctx, h2, weight, targets, ignore_index):
        T, V = h2.shape[0], weight.shape[0]
        ld = (V + 7) // 8 * 8                                    # 16-byte row pitch: TMA operands / vector loads
        buf = torch.empty(T, ld, dtype=torch.bfloat16, device=h2.device)
        logits = buf[:, :V]
        gemm_bf16(h2, weight, out=logits)                        # [T, V] = h W^T
        counted = (targets != ignore_index) & (targets >= 0) & (targets < V)
        inv_n = 1.0 / counted.sum().clamp(min=1).to(torch.float32)
        row_loss = softmax_ce_bf16_(logits, targets, inv_n.reshape(1), ignore_index)
        ctx.save_for_backward(h2, weight, buf)
        ctx.V = V
        return row_loss.sum() * inv_n

    @staticmethod
    def backward(ctx, gout):
        h2, weight, buf = ctx.saved_tensors
        dl = buf[:, :ctx.V]                                       # d loss / d logits, bf16, already scaled by 1 / n
        dh = dw = None
        if ctx.needs_input_grad[0]:
            dh = gemm_bf16(dl, weight, b_mn=True).mul_(gout)      # [T, V] . [V, C]
        if ctx.needs_input_grad[1]:
            dw = gemm_bf16(dl, h2, a_mn=True, b_mn=True, splits=1).mul_(gout)      # dl^T h: [V, C], K = tokens
        return dh, dw, None, None


def lm_head_cross_entropy(h: torch.Tensor, weight: torch.Tensor, targets: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """Mean cross-entropy of ``h @ weight.T`` against ``targets`` (``h`` ``[..., C]``, ``weight`` ``[V, C]``, ``targets`` ``[...]``):
    same value and gradients as ``F.cross_entropy(F.linear(h, weight).float().view(-1, V), targets.view(-1))`` up to bf16
    rounding, on this framework's kernels for bf16 CUDA tensors with ``C % 8 == 0`` and plain PyTorch otherwise. (With every
    target ignored the fused path returns 0 where PyTorch returns NaN.)"""
    V, Cin = weight.shape
    fused = (h.is_cuda and h.dtype == torch.bfloat16 and weight.dtype == torch.bfloat16 and Cin % 8 == 0
             and weight.is_contiguous())
    if not fused:
        return F.cross_entropy(F.linear(h, weight).float().view(-1, V), targets.reshape(-1), ignore_index=ignore_index)
    h2 = h.reshape(-1, Cin)
    if not h2.is_contiguous():
        h2 = h2.contiguous()
    return _LMHeadCEFn.apply(h2, weight, targets.reshape(-1).contiguous(), int(ignore_index))
