"""Modules that run ``Linear [-> ReLU] [-> Dropout]`` as one tcgen05 kernel per direction, under autograd.

``fuse_linear_chains(model)`` rewrites ``nn.Sequential`` containers in place: every
``Linear, ReLU, Dropout`` / ``Linear, ReLU`` / ``Linear`` run is executed by a :class:`_FusedSlot` that sits at the
Linear's index and *shares* its parameters. The Sequential keeps its length and indices (the absorbed
ReLU/Dropout slots become ``nn.Identity``), so ``state_dict`` keys are unchanged
(``linear_relu_stack.0.weight`` ...; ``/root/reference/my_ray_module.py:97-107``).

This is the general path (any model, any optimizer, gradient accumulation). The reference network itself normally
takes the whole-step engine instead (``train.torch.prepare_model`` -> ``parallel.EngineModel``).

Forward (K1-K3): ``y = dropout(relu(x W^T + b))`` in the GEMM epilogue, which also emits ``(y > 0)`` bit-packed.
Backward (K6), for consecutive slots of one Sequential (slot k feeds slot k+1 and nothing else):

* slot k+1's dgrad applies slot k's ReLU/dropout mask in its epilogue from those bits (32 bytes per row and tile instead of
  re-reading the activations) and leaves per-32-row column sums of the masked gradient, so slot k receives its
  pre-activation gradient ready-made and its bias gradient costs a 1/32-size reduction;
* only the LAST slot of a chain masks its incoming gradient itself, with one fused mask+scale+pad kernel
  (``ops.act_backward``);
* ``dW`` runs on the same GEMM with MN-major operands.

In eval mode / without grad the same forward kernel runs with dropout off (K10/K12).
"""
from __future__ import annotations

from typing import Optional

import torch
from torch import nn

from .. import ops

_step_counters = {}


def _next_step(device) -> int:
    k = str(device)
    _step_counters[k] = _step_counters.get(k, 0) + 1
    return _step_counters[k]


class _FusedLinearActFn(torch.autograd.Function):
    """One slot. ``slot`` carries the static chain links: ``prev`` (the slot whose output is this slot's input) and
    ``next`` (the slot that consumes this slot's output)."""

    @staticmethod
    def forward(ctx, x, weight, bias, slot, drop_p, step):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        masked = slot.relu or drop_p > 0
        # the next slot's dgrad will mask with these bits; a chain's last slot masks from y itself
        bits = None
        if masked and slot.next is not None and slot.relu:
            bits = torch.empty(x2.shape[0], (weight.shape[0] + 31) // 32, dtype=torch.int32, device=x2.device)
        y = ops.linear_forward(x2, weight, bias, relu=slot.relu, drop_p=drop_p, seed=slot.seed, step=step, salt=slot.salt,
                               precision=slot.precision, mask_bits_out=bits)
        slot._bits = bits
        prev = slot.prev
        prev_bits = prev._bits if (prev is not None and prev._live_drop_p is not None) else None
        ctx.save_for_backward(x2, weight, y if bits is None and masked else None, prev_bits)
        ctx.slot, ctx.drop_p = slot, drop_p
        ctx.prev_scale = (1.0 / (1.0 - prev._live_drop_p)) if prev_bits is not None and prev._live_drop_p > 0 else 1.0
        ctx.has_bias = bias is not None
        ctx.x_shape = x.shape
        ctx.premasked = bits is not None          # my incoming gradient arrives already masked by the next slot's dgrad
        slot._live_drop_p = drop_p if bits is not None else None
        return y.reshape(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight, y, prev_bits = ctx.saved_tensors
        slot = ctx.slot
        dy2 = dy.reshape(-1, dy.shape[-1])
        if dy2.stride(-1) != 1:
            dy2 = dy2.contiguous()
        n_out = dy2.shape[1]
        n_pad = (n_out + 3) // 4 * 4                # TMA needs a 16-byte row pitch (e.g. the 10-wide logit gradient -> 12)
        db_partial = None
        if ctx.premasked:
            dz = dy2                                 # masked, scaled (and column-summed) by the consumer's dgrad epilogue
            db_partial, slot._db_partial = slot._db_partial, None
            if dz.stride(0) % 4:
                dz = ops.act_backward(dz, None, cols_out=n_pad)
        elif y is not None:
            scale = 1.0 / (1.0 - ctx.drop_p) if ctx.drop_p > 0 else 1.0
            dz = ops.act_backward(dy2, y, scale=scale, cols_out=n_pad)        # mask + scale + pad in one pass
        elif n_pad != n_out or dy2.stride(0) % 4:
            dz = ops.act_backward(dy2, None, cols_out=n_pad)
        else:
            dz = dy2
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            partial = None
            if prev_bits is not None:
                partial = torch.zeros(4 * ((dz.shape[0] + 127) // 128), weight.shape[1], dtype=torch.float32, device=dz.device)
            dx = ops.linear_dgrad(dz, weight, N_out=n_out, mask_bits=prev_bits, mask_scale=ctx.prev_scale,
                                  colsum_partial=partial, precision=slot.precision)
            if slot.prev is not None and prev_bits is not None:
                slot.prev._db_partial = partial
            dx = dx.reshape(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            dw = ops.linear_wgrad(dz, x2, N_out=n_out, precision=slot.precision)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = ops.colsum(db_partial) if db_partial is not None else ops.colsum(dz, cols=n_out)
        return dx, dw, db, None, None, None


def _kernel_ok(x: torch.Tensor, weight: torch.Tensor) -> bool:
    return (x.is_cuda and x.dtype == torch.float32 and weight.dtype == torch.float32
            and weight.shape[1] % 4 == 0 and x.shape[-1] == weight.shape[1])


class _FusedSlot(nn.Module):
    """Sits at the Linear's index inside the Sequential and *is* the Linear's replacement, with identical parameter
    names (``weight`` / ``bias``), so ``state_dict`` keys do not move."""

    def __init__(self, linear: nn.Linear, relu: bool, p_drop: float, salt: int):
        super().__init__()
        self.in_features, self.out_features = linear.in_features, linear.out_features
        self.weight = linear.weight
        self.bias = linear.bias
        self.relu, self.p_drop, self.salt = relu, float(p_drop), salt
        self.seed = 0x5EED
        self.precision = ops.PRECISION_FP32
        # chain links (plain attributes, not submodules): set by fuse_linear_chains for directly connected slots
        object.__setattr__(self, "prev", None)
        object.__setattr__(self, "next", None)
        self._bits: Optional[torch.Tensor] = None          # (y > 0) of the last training forward, bit-packed
        self._db_partial: Optional[torch.Tensor] = None    # column-sum partials of my pre-activation gradient (from next's dgrad)
        self._live_drop_p: Optional[float] = None          # drop probability of that forward (None: no bits emitted)

    def forward(self, x):
        w, b = self.weight, self.bias
        if not _kernel_ok(x, w):
            self._live_drop_p = None
            y = torch.nn.functional.linear(x, w, b)           # CPU workers / unsupported dtypes
            if self.relu:
                y = torch.relu(y)
            return torch.nn.functional.dropout(y, self.p_drop, self.training) if self.p_drop > 0 else y
        p = self.p_drop if self.training else 0.0
        step = _next_step(x.device) if p > 0 else 0
        if torch.is_grad_enabled() and (x.requires_grad or w.requires_grad):
            return _FusedLinearActFn.apply(x, w, b, self, p, step)
        self._live_drop_p = None
        x2 = x.reshape(-1, x.shape[-1])
        y = ops.linear_forward(x2 if x2.is_contiguous() else x2.contiguous(), w, b, relu=self.relu, drop_p=p,
                               seed=self.seed, step=step, salt=self.salt, precision=self.precision)
        return y.reshape(*x.shape[:-1], w.shape[0])

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features}, relu={self.relu}, p_drop={self.p_drop}"


def fuse_linear_chains(model: nn.Module) -> nn.Module:
    """Rewrite ``Linear [, ReLU] [, Dropout]`` runs inside every ``nn.Sequential`` of ``model`` and link directly
    connected slots so that their backward passes share masks and column sums (see the module docstring)."""
    salt = [0]

    def rewrite(seq: nn.Sequential):
        mods = list(seq._modules.items())
        i = 0
        last_slot, last_end = None, -1
        while i < len(mods):
            name, m = mods[i]
            if type(m) is nn.Linear and m.weight.dtype == torch.float32 and m.in_features % 4 == 0:
                relu, p, absorbed = False, 0.0, []
                j = i + 1
                if j < len(mods) and type(mods[j][1]) is nn.ReLU:
                    relu = True
                    absorbed.append(mods[j][0])
                    j += 1
                    if j < len(mods) and type(mods[j][1]) is nn.Dropout:
                        p = mods[j][1].p
                        absorbed.append(mods[j][0])
                        j += 1
                salt[0] += 1
                slot = _FusedSlot(m, relu, p, salt[0])
                seq._modules[name] = slot
                for a in absorbed:
                    seq._modules[a] = nn.Identity()
                # slot k -> slot k+1 with nothing in between, a ReLU on k (its output sign is the mask) and a width the
                # bit-mask epilogues take: k+1's dgrad does k's activation backward
                if last_slot is not None and last_end == i and last_slot.relu and last_slot.out_features % 32 == 0:
                    object.__setattr__(last_slot, "next", slot)
                    object.__setattr__(slot, "prev", last_slot)
                last_slot, last_end = slot, j
                i = j
            else:
                i += 1

    for mod in model.modules():
        if isinstance(mod, nn.Sequential):
            rewrite(mod)
    return model
