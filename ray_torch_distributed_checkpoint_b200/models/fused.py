"""Modules that run ``Linear [-> ReLU] [-> Dropout]`` as one tcgen05 kernel per direction.

``fuse_linear_chains(model)`` rewrites ``nn.Sequential`` containers in place: every
``Linear, ReLU, Dropout`` / ``Linear, ReLU`` / ``Linear`` run is executed by :class:`FusedLinearAct`,
which *shares* the original ``nn.Linear``'s parameters. The Sequential keeps its length and indices
(the absorbed ReLU/Dropout slots become ``nn.Identity``), so ``state_dict`` keys are unchanged
(``linear_relu_stack.0.weight`` ...; ``/root/reference/my_ray_module.py:97-107``).

Forward: ``y = dropout(relu(x W^T + b))`` in the GEMM epilogue (K1-K3). Backward (K6): the ReLU+dropout
mask is recovered from ``y > 0`` (no mask tensor is stored), ``dX`` and ``dW`` run on the same kernel
with MN-major operands, ``db`` is a deterministic column sum. In eval mode / without grad the same
forward kernel runs with dropout off (K10/K12).
"""
from __future__ import annotations

import torch
from torch import nn

from .. import ops

_step_counters = {}


def _next_step(device) -> int:
    k = str(device)
    _step_counters[k] = _step_counters.get(k, 0) + 1
    return _step_counters[k]


class _FusedLinearActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, relu, drop_p, seed, step, salt, precision):
        x2 = x.reshape(-1, x.shape[-1])
        if not x2.is_contiguous():
            x2 = x2.contiguous()
        y = ops.linear_forward(x2, weight, bias, relu=relu, drop_p=drop_p, seed=seed, step=step, salt=salt,
                               precision=precision)
        ctx.save_for_backward(x2, weight, y)
        ctx.relu, ctx.drop_p, ctx.precision = relu, drop_p, precision
        ctx.has_bias = bias is not None
        ctx.x_shape = x.shape
        return y.reshape(*x.shape[:-1], weight.shape[0])

    @staticmethod
    def backward(ctx, dy):
        x2, weight, y = ctx.saved_tensors
        dy2 = dy.reshape(-1, dy.shape[-1])
        if ctx.relu or ctx.drop_p > 0:
            # d(pre-activation) = dy * 1[y > 0] / (1 - p): y is non-zero exactly where the unit was kept
            # by dropout and passed the ReLU
            scale = 1.0 / (1.0 - ctx.drop_p) if ctx.drop_p > 0 else 1.0
            dz = torch.where(y > 0, dy2 * scale, torch.zeros((), device=dy.device, dtype=dy.dtype)) if ctx.relu \
                else (dy2 * scale * (y != 0))
        else:
            dz = dy2
        dz = dz.contiguous()
        n_out = dz.shape[1]
        if n_out % 4 != 0:
            # TMA needs a 16-byte row pitch: pad e.g. the 10-wide logit gradient to 12 columns of zeros
            dz_p = torch.zeros(dz.shape[0], (n_out + 3) // 4 * 4, device=dz.device, dtype=dz.dtype)
            dz_p[:, :n_out] = dz
            dz = dz_p
        dx = dw = db = None
        if ctx.needs_input_grad[0]:
            dx = ops.linear_dgrad(dz, weight, N_out=n_out, precision=ctx.precision).reshape(ctx.x_shape)
        if ctx.needs_input_grad[1]:
            dw = ops.linear_wgrad(dz, x2, N_out=n_out, precision=ctx.precision)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = ops.colsum(dz, cols=n_out)
        return dx, dw, db, None, None, None, None, None, None


def _kernel_ok(x: torch.Tensor, weight: torch.Tensor) -> bool:
    return (x.is_cuda and x.dtype == torch.float32 and weight.dtype == torch.float32
            and weight.shape[1] % 4 == 0 and x.shape[-1] == weight.shape[1])


class FusedLinearAct(nn.Module):
    """``dropout(relu(linear(x)))`` backed by the fused kernels; owns no parameters of its own beyond
    the wrapped ``nn.Linear`` (registered under the name ``linear`` only when constructed standalone)."""

    def __init__(self, linear: nn.Linear, relu: bool = False, p_drop: float = 0.0, seed: int = 0x5EED, salt: int = 0,
                 precision: int = ops.PRECISION_FP32):
        super().__init__()
        self.linear = linear
        self.relu, self.p_drop, self.seed, self.salt, self.precision = relu, float(p_drop), seed, salt, precision

    @property
    def weight(self):
        return self.linear.weight

    @property
    def bias(self):
        return self.linear.bias

    def forward(self, x):
        w, b = self.linear.weight, self.linear.bias
        if not _kernel_ok(x, w):
            y = torch.nn.functional.linear(x, w, b)           # CPU workers / unsupported dtypes
            if self.relu:
                y = torch.relu(y)
            return torch.nn.functional.dropout(y, self.p_drop, self.training) if self.p_drop > 0 else y
        p = self.p_drop if self.training else 0.0
        step = _next_step(x.device) if p > 0 else 0
        if torch.is_grad_enabled() and (x.requires_grad or w.requires_grad):
            return _FusedLinearActFn.apply(x, w, b, self.relu, p, self.seed, step, self.salt, self.precision)
        x2 = x.reshape(-1, x.shape[-1])
        y = ops.linear_forward(x2 if x2.is_contiguous() else x2.contiguous(), w, b, relu=self.relu, drop_p=p,
                               seed=self.seed, step=step, salt=self.salt, precision=self.precision)
        return y.reshape(*x.shape[:-1], w.shape[0])

    def extra_repr(self) -> str:
        return f"in={self.linear.in_features}, out={self.linear.out_features}, relu={self.relu}, p_drop={self.p_drop}"


class _FusedSlot(nn.Module):
    """Sits at the Linear's index inside the Sequential. The original ``nn.Linear`` stays registered
    as ``self`` would shift keys, so this class *is* the Linear's replacement with identical
    parameter names: ``weight`` / ``bias``."""

    def __init__(self, linear: nn.Linear, relu: bool, p_drop: float, salt: int):
        super().__init__()
        self.in_features, self.out_features = linear.in_features, linear.out_features
        self.weight = linear.weight
        self.bias = linear.bias
        self.relu, self.p_drop, self.salt = relu, float(p_drop), salt
        self.seed = 0x5EED
        self.precision = ops.PRECISION_FP32

    def forward(self, x):
        w, b = self.weight, self.bias
        if not _kernel_ok(x, w):
            y = torch.nn.functional.linear(x, w, b)
            if self.relu:
                y = torch.relu(y)
            return torch.nn.functional.dropout(y, self.p_drop, self.training) if self.p_drop > 0 else y
        p = self.p_drop if self.training else 0.0
        step = _next_step(x.device) if p > 0 else 0
        if torch.is_grad_enabled() and (x.requires_grad or w.requires_grad):
            return _FusedLinearActFn.apply(x, w, b, self.relu, p, self.seed, step, self.salt, self.precision)
        x2 = x.reshape(-1, x.shape[-1])
        y = ops.linear_forward(x2 if x2.is_contiguous() else x2.contiguous(), w, b, relu=self.relu, drop_p=p,
                               seed=self.seed, step=step, salt=self.salt, precision=self.precision)
        return y.reshape(*x.shape[:-1], w.shape[0])

    def extra_repr(self) -> str:
        return f"in_features={self.in_features}, out_features={self.out_features}, relu={self.relu}, p_drop={self.p_drop}"


def fuse_linear_chains(model: nn.Module) -> nn.Module:
    """Rewrite ``Linear [, ReLU] [, Dropout]`` runs inside every ``nn.Sequential`` of ``model``."""
    salt = [0]

    def rewrite(seq: nn.Sequential):
        mods = list(seq._modules.items())
        i = 0
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
                seq._modules[name] = _FusedSlot(m, relu, p, salt[0])
                for a in absorbed:
                    seq._modules[a] = nn.Identity()
                i = j
            else:
                i += 1

    for mod in model.modules():
        if isinstance(mod, nn.Sequential):
            rewrite(mod)
    return model
