"""Python face of the tcgen05 GEMM (``csrc/gemm_tf32.cu``): forward, dgrad and wgrad of a Linear
layer with their fused epilogues.

Reference parity: ``nn.Linear``/``nn.ReLU``/``nn.Dropout`` in
``/root/reference/my_ray_module.py:99-106`` and their autograd backward at ``:159``.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import GemmDesc

PRECISION_FP32 = 3   # 3xTF32, fp32-equivalent accuracy (default: the reference trains in fp32)
PRECISION_TF32 = 1

_scratch = {}
_retired = []      # outgrown workspaces: never freed, captured CUDA graphs may still hold their addresses


def scratch_for(tag: str, nbytes: int, device, stream=None) -> torch.Tensor:
    """Zero-initialised per-(device, stream, tag) workspace that only ever grows.

    Kernels leave their arrival counters zero when they finish, so the buffer can be reused by the
    next launch on the same stream without a memset. A workspace that has to grow is REPLACED, not freed: a CUDA graph
    captured earlier keeps replaying kernels that carry the old address (the engine keeps several plans, for different batch
    sizes, on one stream), so the old buffer stays alive for the life of the process. Growing while a stream capture is in
    progress is refused: the allocation would come out of the graph's private pool.
    """
    s = stream if stream is not None else torch.cuda.current_stream(device)
    key = (torch.device(device).index or 0, int(s.cuda_stream), tag)
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nbytes:
        if torch.cuda.is_current_stream_capturing():
            raise RuntimeError(f"workspace '{tag}' would have to be (re)allocated during a CUDA graph capture; run the op once "
                               "eagerly first (the engine's warm-up steps do)")
        if buf is not None:
            _retired.append(buf)
        buf = torch.zeros(max(nbytes, 4096), dtype=torch.uint8, device=device)
        _scratch[key] = buf
    return buf


def _check(t: torch.Tensor, name: str):
    if not t.is_cuda or t.dtype != torch.float32:
        raise ValueError(f"{name} must be a float32 CUDA tensor (got {t.dtype} on {t.device})")
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name} must be 2-D with unit inner stride (got shape {tuple(t.shape)}, strides {t.stride()})")


def gemm(a: torch.Tensor, b: torch.Tensor, out: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         M: Optional[int] = None, N: Optional[int] = None, K: Optional[int] = None,
         bias: Optional[torch.Tensor] = None, relu: bool = False,
         drop_p: float = 0.0, seed: int = 0, step: int = 0, step_tensor: Optional[torch.Tensor] = None, salt: int = 0,
         mask_src: Optional[torch.Tensor] = None, mask_scale: float = 1.0,
         n_store: Optional[int] = None, precision: int = PRECISION_FP32, bn: int = 0, splits: int = 0,
         colsum_partial: Optional[torch.Tensor] = None, a_exact: bool = False, b_exact: bool = False, alpha: float = 1.0,
         row_bias: Optional[torch.Tensor] = None, row_bias_scale: float = 1.0,
         mask_bits_out: Optional[torch.Tensor] = None, mask_bits_in: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """``out[M,N] = epilogue(sum_k A(m,k) B(n,k))`` on the tensor cores.

    ``a`` is ``[M,K]`` (``a_mn=False``) or ``[K,M]`` (``a_mn=True``); ``b`` is ``[N,K]`` or ``[K,N]``.
    Logical sizes default to the tensor shapes; pass ``M/N/K`` to use fewer rows/columns than stored
    (padded buffers). ``a_exact`` / ``b_exact`` promise that the operand's values are exact in tf32 (e.g. raw
    uint8 pixels widened to fp32) so its 3xTF32 correction pass is skipped; the accumulator is mapped as
    ``acc*alpha + bias[n] + row_bias[m]*row_bias_scale`` before the activation epilogue (input-normalisation
    folding, see ``include/rtdc_b200.h``). ``mask_bits_out`` (int32 ``[M, >= ceil(N/32)]``) receives one bit per output
    element (``out > 0``); ``mask_bits_in`` is such a tensor used as the backward mask in place of ``mask_src``.
    """
    _check(a, "a"); _check(b, "b"); _check(out, "out")
    am, ak = (a.shape[1], a.shape[0]) if a_mn else (a.shape[0], a.shape[1])
    bn_, bk = (b.shape[1], b.shape[0]) if b_mn else (b.shape[0], b.shape[1])
    M = am if M is None else M
    N = bn_ if N is None else N
    K = min(ak, bk) if K is None else K
    d = GemmDesc()
    d.a, d.lda, d.a_mn = a.data_ptr(), a.stride(0), int(a_mn)
    d.b, d.ldb, d.b_mn = b.data_ptr(), b.stride(0), int(b_mn)
    d.M, d.N, d.K = M, N, K
    d.out, d.ldo = out.data_ptr(), out.stride(0)
    d.n_store = N if n_store is None else n_store
    d.bias = bias.data_ptr() if bias is not None else None
    d.relu = int(relu)
    d.drop_p = float(drop_p)
    d.seed = seed & 0xFFFFFFFFFFFFFFFF
    d.step_ptr = step_tensor.data_ptr() if step_tensor is not None else None
    d.step_imm = step & 0xFFFFFFFF
    d.salt = salt & 0xFFFFFFFF
    if mask_src is not None:
        _check(mask_src, "mask_src")
        d.mask_src, d.ldm, d.mask_scale = mask_src.data_ptr(), mask_src.stride(0), float(mask_scale)
    d.precision = precision
    d.bn, d.splits = bn, splits
    if colsum_partial is not None:
        if colsum_partial.shape[0] < 4 * ((M + 127) // 128) or colsum_partial.dtype != torch.float32:
            raise ValueError("colsum_partial must be float32 with at least 4*ceil(M/128) rows")
        d.colsum_partial, d.ld_colsum = colsum_partial.data_ptr(), colsum_partial.stride(0)
    d.a_exact, d.b_exact, d.alpha = int(a_exact), int(b_exact), float(alpha)
    if row_bias is not None:
        if row_bias.dtype != torch.float32 or row_bias.numel() < M or not row_bias.is_contiguous():
            raise ValueError("row_bias must be a contiguous float32 tensor with at least M elements")
        d.row_bias, d.row_bias_scale = row_bias.data_ptr(), float(row_bias_scale)
    for name, t in (("mask_bits_out", mask_bits_out), ("mask_bits_in", mask_bits_in)):
        if t is not None:
            if t.dtype != torch.int32 or not t.is_cuda or t.dim() != 2 or t.stride(1) != 1 or t.shape[0] < M or t.shape[1] * 32 < N:
                raise ValueError(f"{name} must be an int32 CUDA tensor of shape [>= M, >= ceil(N/32)] with unit inner stride")
    if mask_bits_out is not None:
        d.mask_bits_out, d.ld_mask_out = mask_bits_out.data_ptr(), mask_bits_out.stride(0)
    if mask_bits_in is not None:
        d.mask_bits_in, d.ld_mask_in, d.mask_scale = mask_bits_in.data_ptr(), mask_bits_in.stride(0), float(mask_scale)
    lib = _lib.load()
    need = lib.rtdc_gemm_scratch_bytes(C.byref(d))
    if need:
        ws = scratch_for("gemm", need, a.device, stream)
        d.scratch, d.scratch_bytes = ws.data_ptr(), ws.numel()
    _lib.check(lib.rtdc_gemm(C.byref(d), _lib.stream_ptr(stream)))
    return out


def gemm_plan(M: int, N: int, K: int, *, b_mn: bool = False, precision: int = PRECISION_FP32, bn: int = 0, splits: int = 0):
    d = GemmDesc()
    d.M, d.N, d.K, d.b_mn, d.precision, d.bn, d.splits = M, N, K, int(b_mn), precision, bn, splits
    out = (C.c_int * 5)()
    _lib.check(_lib.load().rtdc_gemm_plan(C.byref(d), C.byref(out)))
    return dict(bn=out[0], stages=out[1], splits=out[2], tiles=out[3], smem=out[4])


# ------------------------------------------------------------------------------------------------
# Linear-layer shaped helpers

def linear_forward(x, weight, bias=None, *, relu=False, drop_p=0.0, seed=0, step=0, step_tensor=None, salt=0,
                   out=None, n_store=None, precision=PRECISION_FP32, x_exact=False, alpha=1.0, mask_bits_out=None,
                   stream=None):
    """``y = dropout(relu(alpha * (x @ weight.T) + bias))`` in one kernel (K1-K3, K12). ``x_exact``: the values of
    ``x`` are exact in tf32 (raw pixels), see :func:`gemm`."""
    B, N = x.shape[0], weight.shape[0]
    if out is None:
        out = torch.empty(B, n_store or N, dtype=torch.float32, device=x.device)
    return gemm(x, weight, out, M=B, N=N, K=weight.shape[1], bias=bias, relu=relu, drop_p=drop_p, seed=seed, step=step,
                step_tensor=step_tensor, salt=salt, n_store=n_store, precision=precision, a_exact=x_exact, alpha=alpha,
                mask_bits_out=mask_bits_out, stream=stream)


def linear_dgrad(dy, weight, *, N_out=None, mask_src=None, mask_scale=1.0, out=None, precision=PRECISION_FP32,
                 colsum_partial=None, mask_bits=None, stream=None):
    """``dx = (dy @ weight) * (mask_src > 0) * mask_scale`` (K6 dgrad + ReLU/dropout backward of the
    layer below). ``dy`` may be stored wider than ``N_out`` columns (zero padded). ``mask_bits``: the bit-packed
    ``(h > 0)`` a forward GEMM wrote with ``mask_bits_out``, used instead of ``mask_src``."""
    n_out = weight.shape[0] if N_out is None else N_out
    B, K_in = dy.shape[0], weight.shape[1]
    if out is None:
        out = torch.empty(B, K_in, dtype=torch.float32, device=dy.device)
    return gemm(dy, weight, out, b_mn=True, M=B, N=K_in, K=n_out, mask_src=mask_src, mask_scale=mask_scale,
                precision=precision, colsum_partial=colsum_partial, mask_bits_in=mask_bits, stream=stream)


def linear_wgrad(dy, x, *, N_out=None, out=None, precision=PRECISION_FP32, splits=0, x_exact=False, alpha=1.0,
                 row_bias=None, row_bias_scale=1.0, stream=None):
    """``dW[N_out, K_in] = alpha * (dy[:, :N_out].T @ x) + row_bias[:, None] * row_bias_scale`` with split-K over the
    batch (K6 wgrad). ``out`` may be a view into a gradient bucket, so the gradient is produced in place for the
    all-reduce."""
    n_out = dy.shape[1] if N_out is None else N_out
    B, K_in = x.shape
    if out is None:
        out = torch.empty(n_out, K_in, dtype=torch.float32, device=x.device)
    return gemm(dy, x, out, a_mn=True, b_mn=True, M=n_out, N=K_in, K=B, precision=precision, splits=splits,
                b_exact=x_exact, alpha=alpha, row_bias=row_bias, row_bias_scale=row_bias_scale, stream=stream)
