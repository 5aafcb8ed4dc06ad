"""bf16 LayerNorm (forward / backward with the weight and bias gradients in one pass) and bf16 column sums on this
framework's kernels (``csrc/layernorm.cu``).

Not a reference capability (the reference MLP has no normalisation layer); serves the GPT-2 configuration of BASELINE.json,
where the stock LayerNorm kernels and the bias-gradient reductions were ~5 ms of a 24 ms step.
"""
from __future__ import annotations

import ctypes as C

import torch
import torch.nn.functional as F

from . import _lib

MAX_WIDTH = 1024


def _lib_ln():
    lib = _lib.load()
    if not getattr(lib, "_ln_declared", False):
        lib.rtdc_layernorm_fwd_bf16.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_float, C.c_int, C.c_int,
                                                C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p]
        lib.rtdc_layernorm_fwd_bf16.restype = C.c_int
        lib.rtdc_layernorm_bwd_scratch_bytes.argtypes = [C.c_int, C.c_int]
        lib.rtdc_layernorm_bwd_scratch_bytes.restype = C.c_size_t
        lib.rtdc_layernorm_bwd_bf16.argtypes = [C.c_void_p, C.c_longlong, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_int, C.c_int, C.c_void_p, C.c_longlong, C.c_void_p, C.c_void_p, C.c_void_p,
                                                C.c_size_t, C.c_void_p]
        lib.rtdc_layernorm_bwd_bf16.restype = C.c_int
        lib.rtdc_colsum_bf16_scratch_bytes.argtypes = [C.c_int, C.c_int]
        lib.rtdc_colsum_bf16_scratch_bytes.restype = C.c_size_t
        lib.rtdc_colsum_bf16.argtypes = [C.c_void_p, C.c_longlong, C.c_int, C.c_int, C.c_void_p, C.c_int, C.c_void_p, C.c_size_t,
                                         C.c_void_p]
        lib.rtdc_colsum_bf16.restype = C.c_int
        lib._ln_declared = True
    return lib


def colsum_bf16(x: torch.Tensor, out_dtype=torch.bfloat16) -> torch.Tensor:
    """Column sums (fp32 accumulation) of a 2-D bf16 CUDA matrix whose width and row pitch are multiples of 8."""
    if not x.is_cuda or x.dtype != torch.bfloat16 or x.dim() != 2 or x.stride(1) != 1:
        raise ValueError("x must be a 2-D bfloat16 CUDA tensor with unit inner stride")
    rows, cols = x.shape
    lib = _lib_ln()
    out = torch.empty(cols, dtype=out_dtype, device=x.device)
    nbytes = lib.rtdc_colsum_bf16_scratch_bytes(rows, cols)
    scratch = torch.empty(max(1, nbytes // 4), dtype=torch.float32, device=x.device)
    _lib.check(lib.rtdc_colsum_bf16(x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), int(out_dtype == torch.float32),
                                    scratch.data_ptr(), scratch.numel() * 4, _lib.stream_ptr()))
    return out


def _aligned(*tensors) -> bool:
    return all(t is None or t.data_ptr() % 16 == 0 for t in tensors)


class _LayerNormFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, eps):
        Cn = x.shape[-1]
        x2 = x.reshape(-1, Cn)
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        rows = x2.shape[0]
        y = torch.empty_like(x2)
        mean = torch.empty(rows, dtype=torch.float32, device=x.device)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        _lib.check(_lib_ln().rtdc_layernorm_fwd_bf16(x2.data_ptr(), x2.stride(0), weight.data_ptr(), bias.data_ptr() if bias is not None else None,
                                                     float(eps), rows, Cn, y.data_ptr(), y.stride(0), mean.data_ptr(), rstd.data_ptr(),
                                                     _lib.stream_ptr()))
        ctx.save_for_backward(x2, weight, mean, rstd)
        ctx.has_bias = bias is not None
        ctx.x_shape = x.shape
        return y.view(x.shape)

    @staticmethod
    def backward(ctx, dy):
        x2, weight, mean, rstd = ctx.saved_tensors
        rows, Cn = x2.shape
        dy2 = dy.reshape(rows, Cn)
        if not dy2.is_contiguous() or dy2.data_ptr() % 16:
            dy2 = dy2.clone(memory_format=torch.contiguous_format)
        lib = _lib_ln()
        dx = torch.empty_like(x2)
        dw = torch.empty(Cn, dtype=torch.bfloat16, device=x2.device)
        db = torch.empty(Cn, dtype=torch.bfloat16, device=x2.device) if ctx.has_bias else None
        nbytes = lib.rtdc_layernorm_bwd_scratch_bytes(rows, Cn)
        scratch = torch.empty(nbytes // 4, dtype=torch.float32, device=x2.device)
        _lib.check(lib.rtdc_layernorm_bwd_bf16(dy2.data_ptr(), dy2.stride(0), x2.data_ptr(), x2.stride(0), weight.data_ptr(), mean.data_ptr(),
                                               rstd.data_ptr(), rows, Cn, dx.data_ptr(), dx.stride(0), dw.data_ptr(),
                                               db.data_ptr() if db is not None else None, scratch.data_ptr(), scratch.numel() * 4,
                                               _lib.stream_ptr()))
        return dx.view(ctx.x_shape), dw, db, None


def layer_norm_bf16(x: torch.Tensor, weight: torch.Tensor, bias, eps: float = 1e-5) -> torch.Tensor:
    """``F.layer_norm(x, (C,), weight, bias, eps)`` over the last dimension: this framework's kernels for bf16 CUDA tensors with
    ``C % 8 == 0``, ``C <= 1024`` and 16-byte aligned operands, plain PyTorch otherwise."""
    Cn = x.shape[-1]
    fused = (x.is_cuda and x.dtype == torch.bfloat16 and weight is not None and weight.dtype == torch.bfloat16
             and (bias is None or bias.dtype == torch.bfloat16) and Cn % 8 == 0 and Cn <= MAX_WIDTH and x.numel() > 0
             and weight.is_contiguous() and (bias is None or bias.is_contiguous()) and _aligned(x, weight, bias))
    if not fused:
        return F.layer_norm(x, (Cn,), weight, bias, eps)
    return _LayerNormFn.apply(x, weight, bias, eps)


class LayerNormBf16(torch.nn.Module):
    """Drop-in for ``nn.LayerNorm`` over the last dimension (same parameter names, so ``state_dict`` keys do not move)."""

    def __init__(self, ln: torch.nn.LayerNorm):
        super().__init__()
        self.normalized_shape, self.eps = tuple(ln.normalized_shape), ln.eps
        self.weight, self.bias = ln.weight, ln.bias

    def forward(self, x):
        if tuple(x.shape[-1:]) != self.normalized_shape:
            return F.layer_norm(x, self.normalized_shape, self.weight, self.bias, self.eps)      # raises torch's own shape error
        return layer_norm_bf16(x, self.weight, self.bias, self.eps)

    def extra_repr(self) -> str:
        return f"{self.normalized_shape}, eps={self.eps}, kernel=rtdc layernorm bf16"


def use_fused_layernorm(model: torch.nn.Module) -> torch.nn.Module:
    """Replace every affine ``nn.LayerNorm`` over one trailing dimension by :class:`LayerNormBf16`, in place."""
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if type(child) is torch.nn.LayerNorm and len(child.normalized_shape) == 1 and child.elementwise_affine:
                setattr(parent, name, LayerNormBf16(child))
    return model
