"""Optimizers that step a whole flat parameter arena with one fused kernel (K7).

``SGD`` is a drop-in for ``torch.optim.SGD(model.parameters(), lr, momentum=0.9)``
(``/root/reference/my_ray_module.py:142,160``); ``state_dict()`` has exactly torch's layout
(``state[i]['momentum_buffer']``, ``param_groups[0]`` with ``params: [0..n-1]``), so the reference's
checkpoint dict (``optimizer_state_dict``, ``my_ray_module.py:183``) stays interchangeable and an exact
resume can reload it.

Parameters are moved into one contiguous fp32 arena (or the arena of a ``parallel.DistributedDataParallel``
wrapper is adopted as is), so ``step()`` is a single launch reading the gradient arena — which on GPU is the
symmetric buffer the all-reduce has just finished writing. On CPU the same update runs as vectorised
torch ops on the flat arenas.
"""
from __future__ import annotations

from typing import Any, Dict, Iterable, List, Optional

import torch
from torch import nn

from . import ops


def _unwrap(model_or_params):
    """Return (named parameter list in registration order, ddp wrapper or None)."""
    from .parallel.ddp import DistributedDataParallel
    if isinstance(model_or_params, DistributedDataParallel):
        return [p for p in model_or_params.module.parameters() if p.requires_grad], model_or_params
    if isinstance(model_or_params, nn.Module):
        return [p for p in model_or_params.parameters() if p.requires_grad], None
    return [p for p in model_or_params if p.requires_grad], None


class _FlatOptimizer:
    def __init__(self, model_or_params, defaults: Dict[str, Any]):
        params, ddp = _unwrap(model_or_params)
        if not params:
            raise ValueError("optimizer got an empty parameter list")
        self.params: List[torch.nn.Parameter] = params
        self.defaults = defaults
        self.ddp = ddp
        dev = params[0].device
        self.device = dev
        if ddp is not None:
            self.p_arena, self.g_arena = ddp.param_arena, ddp.grad_arena
            self._views = None
        else:
            # private arenas: 128-byte aligned slots in registration order
            align = 32
            offs, cur = [], 0
            for p in params:
                offs.append(cur)
                cur = (cur + p.numel() + align - 1) // align * align
            self.p_arena = torch.zeros(cur, dtype=params[0].dtype, device=dev)
            self.g_arena = torch.zeros(cur, dtype=params[0].dtype, device=dev)
            self._views = []
            with torch.no_grad():
                for p, o in zip(params, offs):
                    v = self.p_arena[o:o + p.numel()].view(p.shape)
                    v.copy_(p.data)
                    p.data = v
                    gv = self.g_arena[o:o + p.numel()].view(p.shape)
                    p.grad = gv
                    self._views.append(gv)
        self._slot = {}
        for i, p in enumerate(params):
            off = (p.data.data_ptr() - self.p_arena.data_ptr()) // self.p_arena.element_size()
            self._slot[i] = (off, p.numel(), tuple(p.shape))
        self.use_kernels = dev.type == "cuda" and self.p_arena.dtype == torch.float32

    # gradients may have been replaced by autograd (zero_grad(set_to_none=True)); pull them back
    def _gather_grads(self):
        if self.ddp is not None:
            return
        for p, gv in zip(self.params, self._views):
            if p.grad is None:
                gv.zero_()
                p.grad = gv
            elif p.grad.data_ptr() != gv.data_ptr():
                gv.copy_(p.grad)
                p.grad = gv

    def zero_grad(self, set_to_none: bool = False):
        """One memset of the gradient arena; gradients stay arena views (``set_to_none`` is ignored)."""
        self.g_arena.zero_()
        if self.ddp is not None:
            self.ddp.zero_grad()
        else:
            for p, gv in zip(self.params, self._views):
                p.grad = gv

    def _view(self, arena: torch.Tensor, i: int) -> torch.Tensor:
        off, n, shape = self._slot[i]
        return arena[off:off + n].view(shape)


class SGD(_FlatOptimizer):
    def __new__(cls, model_or_params, *args, **kwargs):
        # an engine-backed model (train.torch.prepare_model on the reference network) steps through the fused engine
        from .parallel.engine_api import EngineModel, EngineSGD, owner_of
        owner = model_or_params if isinstance(model_or_params, EngineModel) else None
        if owner is None and not isinstance(model_or_params, nn.Module):
            model_or_params = list(model_or_params)
            owner = owner_of(model_or_params)
        if owner is not None:
            return EngineSGD(owner, *args, **kwargs)
        self = super().__new__(cls)
        self._params_arg = model_or_params
        return self

    def __init__(self, model_or_params, lr: float = 1e-3, momentum: float = 0.0, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False):
        super().__init__(self._params_arg, dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay,
                                                nesterov=nesterov))
        self.m_arena = torch.zeros_like(self.p_arena) if momentum != 0 else None
        self._steps = 0
        self.state_dev = ops.SgdState(self.device, lr, momentum, dampening, weight_decay, nesterov) if self.use_kernels else None
        self._lr_dev = lr
        self.param_groups = [dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov,
                                  maximize=False, foreach=None, differentiable=False, fused=None,
                                  params=list(range(len(self.params))))]

    @torch.no_grad()
    def step(self):
        self._gather_grads()
        g = self.param_groups[0]
        if self.use_kernels:
            if g["lr"] != self._lr_dev:                 # LR schedulers write param_groups[0]['lr'] (compare the value as
                self.state_dev.update(lr=g["lr"])       # given: the device mirror holds it rounded to fp32)
                self._lr_dev = g["lr"]
            ops.sgd_step(self.p_arena, self.g_arena, self.m_arena, self.state_dev)
            if self._steps == 0:
                ops.counter_advance(None, self.state_dev)
        else:
            grad = self.g_arena
            if g["weight_decay"]:
                grad = grad.add(self.p_arena, alpha=g["weight_decay"])
            if self.m_arena is not None:
                if self._steps == 0:
                    self.m_arena.copy_(grad)
                else:
                    self.m_arena.mul_(g["momentum"]).add_(grad, alpha=1 - g["dampening"])
                grad = grad.add(self.m_arena, alpha=g["momentum"]) if g["nesterov"] else self.m_arena
            self.p_arena.add_(grad, alpha=-g["lr"])
        self._steps += 1

    def state_dict(self) -> Dict[str, Any]:
        state = {}
        if self.m_arena is not None and self._steps > 0:
            for i in range(len(self.params)):
                state[i] = {"momentum_buffer": self._view(self.m_arena, i)}
        return {"state": state, "param_groups": [dict(self.param_groups[0])]}

    def load_state_dict(self, sd: Dict[str, Any]):
        g = sd["param_groups"][0]
        for k in ("lr", "momentum", "dampening", "weight_decay", "nesterov"):
            self.param_groups[0][k] = g[k]
        if sd["state"]:
            if self.m_arena is None:
                self.m_arena = torch.zeros_like(self.p_arena)
            with torch.no_grad():
                for i, st in sd["state"].items():
                    self._view(self.m_arena, int(i)).copy_(st["momentum_buffer"])
            self._steps = max(self._steps, 1)
        if self.use_kernels:
            self.state_dev.update(lr=g["lr"], momentum=g["momentum"], dampening=g["dampening"],
                                  weight_decay=g["weight_decay"], nesterov=int(g["nesterov"]),
                                  first_step=0 if sd["state"] else 1)
            self._lr_dev = g["lr"]


class Adam(_FlatOptimizer):
    def __init__(self, model_or_params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 0.0,
                 decoupled_weight_decay: bool = False):
        super().__init__(model_or_params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self.decoupled = decoupled_weight_decay
        # bf16 parameters (the GPT-2 / ResNet-50 configurations): moments and a master copy of the parameters in fp32, one fused
        # kernel reads the bf16 gradient arena and rewrites the bf16 parameters from the master
        self.bf16 = self.device.type == "cuda" and self.p_arena.dtype == torch.bfloat16
        if self.bf16:
            self.use_kernels = True
            self.master = self.p_arena.float()
        self.m_arena = torch.zeros_like(self.p_arena, dtype=torch.float32) if self.bf16 else torch.zeros_like(self.p_arena)
        self.v_arena = torch.zeros_like(self.m_arena)
        self._steps = 0
        self.state_dev = ops.AdamState(self.device, lr, betas, eps, weight_decay, decoupled_weight_decay) if self.use_kernels else None
        self._lr_dev = lr
        self.param_groups = [dict(lr=lr, betas=tuple(betas), eps=eps, weight_decay=weight_decay, amsgrad=False, maximize=False,
                                  foreach=None, capturable=False, differentiable=False, fused=None,
                                  decoupled_weight_decay=decoupled_weight_decay, params=list(range(len(self.params))))]

    @torch.no_grad()
    def step(self):
        self._gather_grads()
        g = self.param_groups[0]
        self._steps += 1
        if self.use_kernels:
            if g["lr"] != self._lr_dev:                 # as given, not the fp32-rounded device mirror (that compare never
                self.state_dev.update(lr=g["lr"])       # matched: one device round trip per step)
                self._lr_dev = g["lr"]
            ops.adam_step(self.p_arena, self.g_arena, self.m_arena, self.v_arena, self.state_dev,
                          master=self.master if self.bf16 else None)
            return
        b1, b2 = g["betas"]
        grad = self.g_arena
        if self.decoupled:
            self.p_arena.mul_(1 - g["lr"] * g["weight_decay"])
        elif g["weight_decay"]:
            grad = grad.add(self.p_arena, alpha=g["weight_decay"])
        self.m_arena.mul_(b1).add_(grad, alpha=1 - b1)
        self.v_arena.mul_(b2).addcmul_(grad, grad, value=1 - b2)
        bc1, bc2 = 1 - b1 ** self._steps, 1 - b2 ** self._steps
        denom = (self.v_arena.sqrt() / (bc2 ** 0.5)).add_(g["eps"])
        self.p_arena.addcdiv_(self.m_arena, denom, value=-g["lr"] / bc1)

    def state_dict(self) -> Dict[str, Any]:
        state = {}
        if self._steps > 0:
            for i in range(len(self.params)):
                state[i] = {"step": torch.tensor(float(self._steps)), "exp_avg": self._view(self.m_arena, i),
                            "exp_avg_sq": self._view(self.v_arena, i)}
        return {"state": state, "param_groups": [dict(self.param_groups[0])]}

    def load_state_dict(self, sd: Dict[str, Any]):
        g = sd["param_groups"][0]
        for k in ("lr", "betas", "eps", "weight_decay"):
            self.param_groups[0][k] = g[k]
        with torch.no_grad():
            for i, st in sd["state"].items():
                self._view(self.m_arena, int(i)).copy_(st["exp_avg"])
                self._view(self.v_arena, int(i)).copy_(st["exp_avg_sq"])
                self._steps = int(float(st["step"]))
        if self.use_kernels:
            self.state_dev.update(lr=g["lr"], beta1=g["betas"][0], beta2=g["betas"][1], eps=g["eps"],
                                  weight_decay=g["weight_decay"], step=self._steps)
            self._lr_dev = g["lr"]


class AdamW(Adam):
    def __init__(self, model_or_params, lr: float = 1e-3, betas=(0.9, 0.999), eps: float = 1e-8, weight_decay: float = 1e-2):
        super().__init__(model_or_params, lr, betas, eps, weight_decay, decoupled_weight_decay=True)
