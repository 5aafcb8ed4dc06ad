"""Python face of the bf16 tcgen05 GEMM (``csrc/gemm_bf16.cu``) and a Linear layer built on it.

Not a reference capability (the reference trains a fp32 MLP); serves the bf16 GPT-2 / ResNet-50 north-star
configurations of BASELINE.json, whose ``nn.Linear`` layers otherwise run on cuBLAS.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib

ACT_NONE, ACT_RELU, ACT_GELU_TANH, ACT_GELU_ERF = 0, 1, 2, 3


class GemmBf16Desc(C.Structure):
    _fields_ = [("a", C.c_void_p), ("lda", C.c_int), ("a_mn", C.c_int),
                ("b", C.c_void_p), ("ldb", C.c_int), ("b_mn", C.c_int),
                ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
                ("out", C.c_void_p), ("ldo", C.c_int), ("out_fp32", C.c_int),
                ("bias", C.c_void_p), ("bias_fp32", C.c_int),
                ("act", C.c_int), ("alpha", C.c_float), ("splits", C.c_int)]


def _lib_bf16():
    lib = _lib.load()
    if not getattr(lib, "_bf16_declared", False):
        lib.rtdc_gemm_bf16.argtypes = [C.POINTER(GemmBf16Desc), C.c_void_p]
        lib.rtdc_gemm_bf16.restype = C.c_int
        lib._bf16_declared = True
    return lib


def gemm_bf16(a: torch.Tensor, b: torch.Tensor, out: Optional[torch.Tensor] = None, *, a_mn: bool = False, b_mn: bool = False,
              bias: Optional[torch.Tensor] = None, act: int = ACT_NONE, alpha: float = 1.0, out_dtype=torch.bfloat16,
              splits: int = 0, stream=None) -> torch.Tensor:
    """``out[M,N] = act(alpha * sum_k A(m,k) B(n,k) + bias[n])`` with bf16 operands and fp32 accumulation.

    ``a`` is ``[M,K]`` (``a_mn=False``) or ``[K,M]`` (``a_mn=True``); ``b`` is ``[N,K]`` or ``[K,N]``; both row-major bf16 with a
    row pitch that is a multiple of 8 elements. ``out`` is bf16 or fp32 (``out_dtype`` when allocated here). ``splits`` > 1
    (fp32 output only, no bias / activation; 0 lets the library choose) splits K across CTAs that add with fp32 atomics."""
    for name, t in (("a", a), ("b", b)):
        if not t.is_cuda or t.dtype != torch.bfloat16 or t.dim() != 2 or t.stride(1) != 1:
            raise ValueError(f"{name} must be a 2-D bfloat16 CUDA tensor with unit inner stride")
    M, Ka = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    N, Kb = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    if Ka != Kb:
        raise ValueError(f"inner dimensions differ: {Ka} vs {Kb}")
    if out is None:
        out = torch.empty(M, N, dtype=out_dtype, device=a.device)
    if out.dtype not in (torch.bfloat16, torch.float32) or out.dim() != 2 or out.stride(1) != 1 or out.shape[0] < M or out.shape[1] < N:
        raise ValueError("out must be a 2-D bf16 / fp32 tensor of at least [M, N] with unit inner stride")
    d = GemmBf16Desc()
    d.a, d.lda, d.a_mn = a.data_ptr(), a.stride(0), int(a_mn)
    d.b, d.ldb, d.b_mn = b.data_ptr(), b.stride(0), int(b_mn)
    d.M, d.N, d.K = M, N, Ka
    d.out, d.ldo, d.out_fp32 = out.data_ptr(), out.stride(0), int(out.dtype == torch.float32)
    if bias is not None:
        if bias.dtype not in (torch.bfloat16, torch.float32) or bias.numel() < N or not bias.is_contiguous():
            raise ValueError("bias must be a contiguous bf16 / fp32 vector of at least N elements")
        d.bias, d.bias_fp32 = bias.data_ptr(), int(bias.dtype == torch.float32)
    d.act, d.alpha, d.splits = int(act), float(alpha), int(splits)
    _lib.check(_lib_bf16().rtdc_gemm_bf16(C.byref(d), _lib.stream_ptr(stream)))
    return out


class _LinearBf16Fn(torch.autograd.Function):
    """``y = x W^T + b`` with forward, dgrad and wgrad on the bf16 tcgen05 GEMM (operands read as stored)."""

    @staticmethod
    def forward(ctx, x, weight, bias):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        y = gemm_bf16(x2, weight, bias=bias)
        ctx.save_for_backward(x2, weight)
        ctx.has_bias = bias is not None
        ctx.x_shape = x.shape
        return y.reshape(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if not dy2.is_contiguous():
            dy2 = dy2.contiguous()
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = gemm_bf16(dy2, weight, b_mn=True).reshape(ctx.x_shape)                     # dy [M, N_out] . W [N_out, K_in]
        if ctx.needs_input_grad[1]:
            dw32 = gemm_bf16(dy2, x2, a_mn=True, b_mn=True, out_dtype=torch.float32)      # dy^T x, K = tokens (split, fp32)
            dw = dw32.to(torch.bfloat16)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            if dy2.shape[1] % 8 == 0 and dy2.data_ptr() % 16 == 0:
                from .layernorm import colsum_bf16
                db = colsum_bf16(dy2)                                                       # csrc/layernorm.cu, fp32 accumulation
            else:
                db = dy2.sum(0, dtype=torch.float32).to(torch.bfloat16)
        return dx, dw, db


class LinearBf16(torch.nn.Module):
    """Drop-in for a bf16 ``nn.Linear`` (same parameter names, so ``state_dict`` keys do not move)."""

    def __init__(self, linear: torch.nn.Linear):
        super().__init__()
        self.in_features, self.out_features = linear.in_features, linear.out_features
        self.weight, self.bias = linear.weight, linear.bias

    def forward(self, x):
        w = self.weight
        ok = (x.is_cuda and x.dtype == torch.bfloat16 and w.dtype == torch.bfloat16 and self.in_features % 8 == 0
              and self.out_features % 8 == 0)
        if not ok:
            return torch.nn.functional.linear(x, w, self.bias)
        if torch.is_grad_enabled() and (x.requires_grad or w.requires_grad):
            return _LinearBf16Fn.apply(x, w, self.bias)
        x2 = x.reshape(-1, x.shape[-1])
        return gemm_bf16(x2 if x2.is_contiguous() else x2.contiguous(), w, bias=self.bias).reshape(*x.shape[:-1], self.out_features)

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features}, bias={self.bias is not None}, kernel=tcgen05 bf16"


def use_bf16_linears(model: torch.nn.Module) -> torch.nn.Module:
    """Replace every bf16 ``nn.Linear`` of ``model`` (sizes multiples of 8) by :class:`LinearBf16`, in place."""
    for parent in list(model.modules()):
        for name, child in list(parent.named_children()):
            if type(child) is torch.nn.Linear and child.weight.dtype == torch.bfloat16 and child.in_features % 8 == 0 \
                    and child.out_features % 8 == 0:
                setattr(parent, name, LinearBf16(child))
    return model
