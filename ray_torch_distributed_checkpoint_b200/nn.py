"""Drop-in loss modules backed by the fused kernels.

``CrossEntropyLoss`` replaces ``torch.nn.CrossEntropyLoss()`` as used at
``/root/reference/my_ray_module.py:141,156``: mean reduction over the batch, class-index targets. On a
CUDA fp32 ``[B, C<=32]`` input the forward kernel computes the loss *and* ``dlogits`` in one pass
(K4/K5), so backward is a scale by the incoming gradient; anything else falls back to
``torch.nn.functional.cross_entropy``.
"""
from __future__ import annotations

import torch
from torch import nn

from . import ops


class _FusedCE(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, target):
        z = logits if logits.is_contiguous() else logits.contiguous()
        need_grad = logits.requires_grad
        dz = torch.empty_like(z) if need_grad else None
        loss = ops.softmax_cross_entropy(z, target, dlogits=dz)
        ctx.save_for_backward(dz) if need_grad else None
        ctx.has_grad = need_grad
        return loss

    @staticmethod
    def backward(ctx, grad_out):
        if not ctx.has_grad:
            return None, None
        (dz,) = ctx.saved_tensors
        return dz * grad_out, None


def __getattr__(name):
    if name == "EvalAccumulator":
        from .parallel.engine_api import EvalAccumulator
        return EvalAccumulator
    raise AttributeError(name)


class CrossEntropyLoss(nn.Module):
    def __init__(self, reduction: str = "mean"):
        super().__init__()
        if reduction != "mean":
            raise ValueError("the fused CrossEntropyLoss implements reduction='mean' (the reference's default)")
        self.reduction = reduction

    def forward(self, logits: torch.Tensor, target: torch.Tensor) -> torch.Tensor:
        from .parallel.engine_api import LazyLogits
        if isinstance(logits, LazyLogits):
            # engine-backed model: hand the batch to the fused step (training) or run the fused eval graph (eval mode)
            return logits.owner._loss(logits, target)
        if (logits.is_cuda and logits.dtype == torch.float32 and logits.dim() == 2 and logits.shape[1] <= 32
                and target.dtype == torch.int64 and target.dim() == 1):
            return _FusedCE.apply(logits, target.contiguous())
        return torch.nn.functional.cross_entropy(logits, target)
