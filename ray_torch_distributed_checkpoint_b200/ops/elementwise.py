"""Python face of ``csrc/elementwise.cu``: fused softmax-cross-entropy, column sums, uint8 input
normalisation, fused SGD / Adam steps over flat arenas.

Reference parity: ``nn.CrossEntropyLoss`` (``/root/reference/my_ray_module.py:141,156,159``),
``torch.optim.SGD(lr, momentum=0.9)`` (``:142,160``), ``ToTensor()+Normalize((0.5,),(0.5,))`` (``:38``),
validation accuracy count (``:170``).
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch

from . import _lib
from ._lib import AdamHyper, SgdHyper
from .gemm import scratch_for


def softmax_cross_entropy(logits: torch.Tensor, labels: torch.Tensor, *, classes: Optional[int] = None,
                          dlogits: Optional[torch.Tensor] = None, grad_scale: Optional[float] = None,
                          relu_mask: bool = False, loss_out: Optional[torch.Tensor] = None,
                          correct_accum: Optional[torch.Tensor] = None, loss_sum_accum: Optional[torch.Tensor] = None,
                          dbias_out: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """Mean cross-entropy of ``logits[:, :classes]``; optionally also writes
    ``dlogits = (softmax - onehot) * grad_scale`` (default ``1/rows``), masked by ``logits > 0`` when
    ``relu_mask`` (the reference model ends in a ReLU), and accumulates the number of correct
    argmax predictions into ``correct_accum`` (uint64/int64 scalar on the device). ``dbias_out[classes]``
    receives the column sums of ``dlogits`` (bias gradient of the logit layer) from the same pass."""
    if logits.dtype != torch.float32 or not logits.is_cuda or logits.dim() != 2 or logits.stride(1) != 1:
        raise ValueError("logits must be a 2-D float32 CUDA tensor with unit inner stride")
    if labels.dtype != torch.int64 or not labels.is_contiguous():
        raise ValueError("labels must be a contiguous int64 tensor")
    rows = logits.shape[0]
    classes = logits.shape[1] if classes is None else classes
    if loss_out is None:
        loss_out = torch.empty((), dtype=torch.float32, device=logits.device)
    if dlogits is not None and (dlogits.stride(0) != logits.stride(0) or dlogits.dtype != torch.float32):
        raise ValueError("dlogits must match the logits layout")
    lib = _lib.load()
    ws = scratch_for("ce", lib.rtdc_softmax_ce_scratch_bytes(rows), logits.device, stream)
    gs = (1.0 / rows) if grad_scale is None else grad_scale
    _lib.check(lib.rtdc_softmax_ce(
        logits.data_ptr(), logits.stride(0), labels.data_ptr(), rows, classes, loss_out.data_ptr(),
        dlogits.data_ptr() if dlogits is not None else None, gs, int(relu_mask),
        correct_accum.data_ptr() if correct_accum is not None else None,
        loss_sum_accum.data_ptr() if loss_sum_accum is not None else None,
        dbias_out.data_ptr() if dbias_out is not None else None,
        ws.data_ptr(), ws.numel(), _lib.stream_ptr(stream)))
    return loss_out


def mlp_head_train(h: torch.Tensor, weight: torch.Tensor, bias: Optional[torch.Tensor], labels: torch.Tensor, *,
                   dh: torch.Tensor, dw: torch.Tensor, db: torch.Tensor, loss: Optional[torch.Tensor] = None,
                   grad_scale: Optional[float] = None, mask_scale: float = 1.0, db_below: Optional[torch.Tensor] = None,
                   stream=None, part: int = 0, scratch_stream=None) -> torch.Tensor:
    """Fused last layer for training: ``z = relu(h @ weight.T + bias)``, mean cross-entropy, and the layer's
    backward (``dw``, ``db``, ``dh = (dz @ weight) * (h > 0) * mask_scale``) in one pass over ``h``.
    ``weight`` is ``[classes, K]`` with 10 or 16 classes and ``K <= 512``. ``db_below[K]`` optionally receives the
    column sums of ``dh`` (the bias gradient of the layer that produced ``h``). ``part`` 1 / 2 run only the pass over
    ``h`` / only the reduction of its per-block partials (same arguments; ``scratch_stream`` names the stream whose scratch
    buffer both calls share when the reduction goes to a side stream)."""
    rows, K = h.shape
    classes = weight.shape[0]
    if loss is None:
        loss = torch.empty((), dtype=torch.float32, device=h.device)
    lib = _lib.load()
    ws = scratch_for("head", lib.rtdc_mlp_head_scratch_bytes(rows, K, classes), h.device,
                     scratch_stream if scratch_stream is not None else stream)
    gs = (1.0 / rows) if grad_scale is None else grad_scale
    _lib.check(lib.rtdc_mlp_head_train_part(h.data_ptr(), h.stride(0), weight.data_ptr(), bias.data_ptr() if bias is not None else None,
                                            labels.data_ptr(), rows, K, classes, gs, mask_scale, dh.data_ptr(), dh.stride(0),
                                            dw.data_ptr(), db.data_ptr(), loss.data_ptr(),
                                            db_below.data_ptr() if db_below is not None else None, ws.data_ptr(), ws.numel(),
                                            _lib.stream_ptr(stream), part))
    return loss


def colsum(x: torch.Tensor, *, cols: Optional[int] = None, out: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """``out[c] = sum_r x[r, c]`` for the first ``cols`` columns (bias gradient)."""
    if x.dtype != torch.float32 or not x.is_cuda or x.dim() != 2 or x.stride(1) != 1:
        raise ValueError("x must be a 2-D float32 CUDA tensor with unit inner stride")
    rows = x.shape[0]
    cols = x.shape[1] if cols is None else cols
    if out is None:
        out = torch.empty(cols, dtype=torch.float32, device=x.device)
    lib = _lib.load()
    ws = scratch_for("colsum", lib.rtdc_colsum_scratch_bytes(rows, cols), x.device, stream)
    _lib.check(lib.rtdc_colsum(x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), ws.data_ptr(), ws.numel(),
                               _lib.stream_ptr(stream)))
    return out


def act_backward(dy: torch.Tensor, y: Optional[torch.Tensor], *, scale: float = 1.0, cols_out: Optional[int] = None,
                 out: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """``dz = dy * (y > 0) * scale`` (ReLU / inverted-dropout backward; ``y`` None: plain copy), written ``cols_out`` wide with
    zero padding so the row pitch suits the TMA-fed GEMMs that consume it."""
    if dy.dtype != torch.float32 or not dy.is_cuda or dy.dim() != 2 or dy.stride(1) != 1:
        raise ValueError("dy must be a 2-D float32 CUDA tensor with unit inner stride")
    if y is not None and (y.dtype != torch.float32 or y.shape != dy.shape or y.stride(1) != 1):
        raise ValueError("y must match dy")
    rows, cols = dy.shape
    cols_out = cols if cols_out is None else cols_out
    if out is None:
        out = torch.empty(rows, cols_out, dtype=torch.float32, device=dy.device)
    _lib.check(_lib.load().rtdc_act_backward(dy.data_ptr(), dy.stride(0), y.data_ptr() if y is not None else None,
                                             y.stride(0) if y is not None else 0, rows, cols, cols_out, float(scale),
                                             out.data_ptr(), out.stride(0), _lib.stream_ptr(stream)))
    return out


def u8_gather_rows(src: torch.Tensor, idx: torch.Tensor, out: torch.Tensor, *, labels: Optional[torch.Tensor] = None,
                   y_out: Optional[torch.Tensor] = None, stream=None) -> torch.Tensor:
    """``out[r] = src[idx[r]]`` over a device-resident uint8 matrix (row length a multiple of 16 bytes), and
    ``y_out[r] = labels[idx[r]]`` widened to int64: a shuffled batch assembled at HBM speed."""
    if src.dtype != torch.uint8 or out.dtype != torch.uint8 or idx.dtype != torch.int32 or not (src.is_cuda and out.is_cuda and idx.is_cuda):
        raise ValueError("src/out must be uint8 and idx int32 CUDA tensors")
    rows = idx.numel()
    if out.shape[0] < rows or not (src.is_contiguous() and out.is_contiguous() and idx.is_contiguous()):
        raise ValueError("out too small or tensors not contiguous")
    if labels is not None and (labels.dtype != torch.uint8 or y_out is None or y_out.dtype != torch.int64 or y_out.numel() < rows):
        raise ValueError("labels must be uint8 with an int64 y_out of at least idx.numel() elements")
    _lib.check(_lib.load().rtdc_u8_gather_rows(src.data_ptr(), idx.data_ptr(), rows, src.shape[1], out.data_ptr(),
                                               labels.data_ptr() if labels is not None else None,
                                               y_out.data_ptr() if labels is not None else None, _lib.stream_ptr(stream)))
    return out


def h2d_pieces(dst: torch.Tensor, src_host: torch.Tensor, offsets: torch.Tensor, sizes: torch.Tensor, stream=None) -> None:
    """Copy pieces of the pinned host tensor ``src_host`` (byte ``offsets`` / ``sizes``: int64 CPU tensors) back to back into the
    device tensor ``dst`` with one ``cudaMemcpyAsync`` each, issued from native code."""
    if not src_host.is_pinned() or not dst.is_cuda or offsets.dtype != torch.int64 or sizes.dtype != torch.int64 or offsets.is_cuda:
        raise ValueError("src_host must be pinned, dst on the device, offsets/sizes int64 CPU tensors")
    if int(sizes.sum()) > dst.numel() * dst.element_size():
        raise ValueError("pieces do not fit the destination")
    _lib.check(_lib.load().rtdc_h2d_pieces(dst.data_ptr(), src_host.data_ptr(), offsets.data_ptr(), sizes.data_ptr(), offsets.numel(),
                                           _lib.stream_ptr(stream)))


def u8_normalize(src: torch.Tensor, *, mean: float = 0.5, std: float = 0.5, out: Optional[torch.Tensor] = None,
                 stream=None) -> torch.Tensor:
    """uint8 pixels -> float32 ``(x/255 - mean)/std``, any shape (contiguous)."""
    if src.dtype != torch.uint8 or not src.is_cuda or not src.is_contiguous():
        raise ValueError("src must be a contiguous uint8 CUDA tensor")
    if out is None:
        out = torch.empty(src.shape, dtype=torch.float32, device=src.device)
    _lib.check(_lib.load().rtdc_u8_normalize(src.data_ptr(), out.data_ptr(), src.numel(), mean, std, _lib.stream_ptr(stream)))
    return out


def u8_affine(src: torch.Tensor, *, scale: float = 1.0, shift: float = 0.0, out: Optional[torch.Tensor] = None,
              blocks: int = 0, stream=None) -> torch.Tensor:
    """uint8 -> float32 ``x*scale + shift``; the defaults widen raw pixels exactly (every value is then exact in
    tf32, which is what ``gemm(..., a_exact/b_exact)`` wants). ``blocks`` caps the grid (0 = full speed) for a
    conversion that shares the GPU with a tensor-bound kernel."""
    if src.dtype != torch.uint8 or not src.is_cuda or not src.is_contiguous():
        raise ValueError("src must be a contiguous uint8 CUDA tensor")
    if out is None:
        out = torch.empty(src.shape, dtype=torch.float32, device=src.device)
    _lib.check(_lib.load().rtdc_u8_affine_paced(src.data_ptr(), out.data_ptr(), src.numel(), scale, shift, blocks,
                                                _lib.stream_ptr(stream)))
    return out


def fold_input_bias(weight: torch.Tensor, bias: Optional[torch.Tensor], t: float, *, out: Optional[torch.Tensor] = None,
                    stream=None) -> torch.Tensor:
    """``out[n] = bias[n] + t * weight[n, :].sum()``: the bias a Linear layer sees once the input normalisation
    ``x = s*raw + t`` is folded into its GEMM (``W x + b = s*(W raw) + out``)."""
    if weight.dtype != torch.float32 or not weight.is_cuda or weight.dim() != 2 or weight.stride(1) != 1:
        raise ValueError("weight must be a 2-D float32 CUDA tensor with unit inner stride")
    rows, cols = weight.shape
    if out is None:
        out = torch.empty(rows, dtype=torch.float32, device=weight.device)
    _lib.check(_lib.load().rtdc_fold_input_bias(weight.data_ptr(), weight.stride(0), rows, cols,
                                                bias.data_ptr() if bias is not None else None, t, out.data_ptr(),
                                                _lib.stream_ptr(stream)))
    return out


def _hyper_tensor(struct, device) -> torch.Tensor:
    raw = bytes(struct)
    return torch.frombuffer(bytearray(raw), dtype=torch.uint8).to(device)


class SgdState:
    """Device-resident SGD hyper-parameters (readable by captured CUDA graphs) + host mirror."""

    def __init__(self, device, lr, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, grad_scale=1.0):
        self.host = SgdHyper(lr, momentum, dampening, weight_decay, grad_scale, int(nesterov), 1, 0)
        self.dev = _hyper_tensor(self.host, device)

    def update(self, **kw):
        for k, v in kw.items():
            setattr(self.host, k, v)
        self.dev.copy_(torch.frombuffer(bytearray(bytes(self.host)), dtype=torch.uint8), non_blocking=False)


def sgd_step(p: torch.Tensor, g: torch.Tensor, m: Optional[torch.Tensor], state: SgdState, stream=None, *,
             advance: bool = False, counter: Optional[torch.Tensor] = None, loss: Optional[torch.Tensor] = None,
             loss_log: Optional[torch.Tensor] = None):
    """One fused SGD(+momentum) step over flat fp32 arenas (in place on ``p`` and ``m``). ``advance``: the kernel also
    clears ``first_step`` and adds 1 to ``counter`` (what :func:`counter_advance` does in its own launch); with ``loss``
    (device scalar) and ``loss_log`` (pinned host ring, power-of-two length) it also stores the step's loss into
    ``loss_log[counter % len]`` straight over PCIe."""
    n = p.numel()
    if g.numel() != n or (m is not None and m.numel() != n):
        raise ValueError("arena sizes differ")
    lib = _lib.load()
    mp = m.data_ptr() if m is not None else p.data_ptr()
    if advance:
        if loss_log is not None and not loss_log.is_pinned():
            raise ValueError("loss_log must be pinned host memory")
        _lib.check(lib.rtdc_sgd_step_advance_log(p.data_ptr(), g.data_ptr(), mp, n, state.dev.data_ptr(),
                                                 counter.data_ptr() if counter is not None else None,
                                                 loss.data_ptr() if loss is not None else None,
                                                 loss_log.data_ptr() if loss_log is not None else None,
                                                 loss_log.numel() if loss_log is not None else 0, _lib.stream_ptr(stream)))
        state.host.first_step = 0
    else:
        _lib.check(lib.rtdc_sgd_step(p.data_ptr(), g.data_ptr(), mp, n, state.dev.data_ptr(), _lib.stream_ptr(stream)))


def counter_advance(counter: Optional[torch.Tensor], sgd_state: Optional[SgdState] = None, stream=None):
    """``counter += 1`` on the stream; also clears ``first_step`` of an SGD state."""
    _lib.check(_lib.load().rtdc_counter_advance(counter.data_ptr() if counter is not None else None,
                                                sgd_state.dev.data_ptr() if sgd_state is not None else None,
                                                _lib.stream_ptr(stream)))
    if sgd_state is not None:
        sgd_state.host.first_step = 0


class AdamState:
    def __init__(self, device, lr, betas=(0.9, 0.999), eps=1e-8, weight_decay=0.0, decoupled=False, grad_scale=1.0):
        self.host = AdamHyper(lr, betas[0], betas[1], eps, weight_decay, grad_scale, int(decoupled), 0)
        self.dev = _hyper_tensor(self.host, device)

    def update(self, **kw):
        # the step counter lives on the device; preserve it
        cur = AdamHyper.from_buffer_copy(bytes(self.dev.cpu().numpy().tobytes()))
        for k, v in kw.items():
            setattr(cur, k, v)
        self.host = cur
        self.dev.copy_(torch.frombuffer(bytearray(bytes(cur)), dtype=torch.uint8))


def adam_step(p, g, m, v, state: AdamState, stream=None, master=None):
    """Advance the device step counter, then one fused Adam/AdamW step over flat arenas: fp32 ``p`` / ``g``, or bf16 ``p`` /
    ``g`` with an fp32 ``master`` copy of the parameters (``m`` / ``v`` are fp32 either way)."""
    lib = _lib.load()
    sp = _lib.stream_ptr(stream)
    _lib.check(lib.rtdc_adam_advance(state.dev.data_ptr(), sp))
    if p.dtype == torch.bfloat16:
        if master is None or master.dtype != torch.float32 or g.dtype != torch.bfloat16:
            raise ValueError("bf16 parameters need bf16 gradients and an fp32 master copy")
        _lib.check(lib.rtdc_adam_step_bf16(p.data_ptr(), g.data_ptr(), master.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(),
                                           state.dev.data_ptr(), sp))
        return
    _lib.check(lib.rtdc_adam_step(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), p.numel(), state.dev.data_ptr(), sp))
