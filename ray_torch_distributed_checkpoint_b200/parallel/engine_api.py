"""The reference's literal training loop on the fused engine.

``train.torch.prepare_model`` returns an :class:`EngineModel` when it recognises the reference network on a GPU
worker. With it, the loop of ``/root/reference/my_ray_module.py:154-160`` ::

    pred = model(X)
    loss = loss_fn(pred, y)          # ray_torch_distributed_checkpoint_b200.nn.CrossEntropyLoss
    optimizer.zero_grad()
    loss.backward()
    optimizer.step()                 # ray_torch_distributed_checkpoint_b200.optim.SGD

runs as ONE replayed CUDA graph per step (``parallel.engine.MLPTrainEngine``): ``model(X)`` and ``loss_fn`` only
record what was asked for (:class:`LazyLogits`, :class:`LazyLoss`) and hand the batch to the engine's staging stream,
``loss.backward()`` marks the loss as wanted, and ``optimizer.step()`` fires forward + loss + backward + gradient
reduction + SGD. Nothing is computed twice and nothing is computed that was not asked for:

* ``loss.item()`` after ``optimizer.step()`` reads the loss the step computed (from the pinned loss ring);
  before ``step()`` it costs one extra training-mode forward with the same dropout masks;
* in eval mode (``model.eval()`` / ``torch.no_grad()``) ``loss_fn(pred, y)`` runs the fused forward + loss graph at
  once and also accumulates the loss sum and the number of correct predictions on the device
  (:class:`EvalAccumulator` reads them with one D2H per epoch; the reference does two per batch, ``:168,170``);
* ``pred`` turns into a real ``[B, classes]`` tensor on first use as one (``pred.argmax(1)``, ``torch.as_tensor``...).

Batches of any size work: the model keeps one engine plan per batch size it meets (full batch, the epoch's tail batch,
the validation batch). What the engine cannot express (two forwards before one ``step()``, gradient accumulation,
a foreign optimizer) raises with a pointer to ``prepare_model(model, fuse="modules")``, the general autograd path.
"""
from __future__ import annotations

import weakref
from typing import Any, Dict, Iterable, List, Optional

import torch
from torch import nn

from .engine import LOSS_LOG_SLOTS, EngineCore, match_reference_mlp

_OWNER_OF_PARAM: "weakref.WeakValueDictionary[int, EngineModel]" = weakref.WeakValueDictionary()


def owner_of(params: Iterable[torch.nn.Parameter]) -> Optional["EngineModel"]:
    """The EngineModel whose parameters these are (all of them), else None."""
    owner = None
    n = 0
    for p in params:
        o = _OWNER_OF_PARAM.get(id(p))
        if o is None or (owner is not None and o is not owner):
            return None
        owner = o
        n += 1
    if owner is None or n != len(owner.core.param_names):
        return None
    return owner


def _raw_u8(x):
    from ..data.loader import U8Images
    if isinstance(x, U8Images):
        return x.raw
    return x


class LazyLogits:
    """``model(X)`` on the engine path: the logits of one batch, computed on first use."""

    def __init__(self, owner: "EngineModel", x: torch.Tensor, training: bool):
        self.owner, self.x, self.training = owner, x, training
        self.batch = int(x.shape[0])
        self._value: Optional[torch.Tensor] = None
        self._eval_seq: Optional[int] = None       # owner._eval_seq right after the eval graph that ran on this batch
        self._loss: Optional["LazyLoss"] = None

    @property
    def shape(self):
        return torch.Size((self.batch, self.owner.core.d_out))

    @property
    def device(self):
        return self.owner.core.device

    @property
    def dtype(self):
        return torch.float32

    def materialize(self) -> torch.Tensor:
        if self._value is not None:
            return self._value
        o = self.owner
        plan = o.core.plan(self.batch)
        if self.training:
            if self._loss is not None and self._loss.state == "stepped":
                raise RuntimeError("training-mode logits were not kept: the fused head never writes them. Read them before "
                                   "optimizer.step(), or use prepare_model(model, fuse='modules')")
            if self._loss is not None and self._loss.state == "staged":
                plan.forward_loss()                            # the batch is staged already; it stays staged for step()
                self._value = plan.z[:, :o.core.d_out].clone()
            else:
                self._value = plan.predict(_raw_u8(self.x), train=True)
        elif self._eval_seq is not None:
            if self._eval_seq != o._eval_seq:
                raise RuntimeError("these logits were overwritten by a later forward pass; use them before the next batch")
            self._value = plan.z[:, :o.core.d_out].clone()
        else:
            self._value = plan.predict(_raw_u8(self.x))
        return self._value

    # the tensor methods evaluation code reaches for; anything else goes through __getattr__
    def argmax(self, *a, **k):
        return self.materialize().argmax(*a, **k)

    def __getattr__(self, name):
        if name.startswith("__"):
            raise AttributeError(name)
        return getattr(self.materialize(), name)

    def __array__(self, *a, **k):
        return self.materialize().cpu().numpy()

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        conv = lambda v: v.materialize() if isinstance(v, LazyLogits) else v
        return func(*[conv(a) for a in args], **{k: conv(v) for k, v in (kwargs or {}).items()})

    def __repr__(self):
        return f"LazyLogits(batch={self.batch}, classes={self.owner.core.d_out}, training={self.training}, materialized={self._value is not None})"


class LazyLoss:
    """``loss_fn(pred, y)`` on the engine path in training mode. States: ``staged`` (batch handed to the engine),
    ``wanted`` (``backward()`` called), ``stepped`` (``optimizer.step()`` ran the fused step)."""

    def __init__(self, owner: "EngineModel", logits: LazyLogits, plan):
        self.owner, self.logits, self.plan = owner, logits, plan
        self.state = "staged"
        self.step_index: Optional[int] = None
        self._peek: Optional[float] = None
        self.requires_grad = True

    def backward(self, *a, **k):
        if self.owner._optimizer is None:
            raise RuntimeError("the engine-backed model needs ray_torch_distributed_checkpoint_b200.optim.SGD as its optimizer "
                               "(a foreign optimizer never sees gradients here); use prepare_model(model, fuse='modules') for "
                               "the general autograd path")
        if self.state == "stepped":
            raise RuntimeError("backward() after optimizer.step(): this loss has been consumed")
        self.state = "wanted"

    def item(self) -> float:
        core = self.owner.core
        if self.state == "stepped":
            if core.log_index - self.step_index > LOSS_LOG_SLOTS:
                raise RuntimeError("this loss is older than the loss ring")
            core.stream.synchronize()
            return float(core.loss_log[self.step_index % LOSS_LOG_SLOTS])
        if self._peek is None:
            self._peek = float(self.plan.forward_loss().item())
        return self._peek

    __float__ = item

    def detach(self):
        return torch.tensor(self.item())

    def __repr__(self):
        return f"LazyLoss(state={self.state})"


class EngineModel(nn.Module):
    """The reference network backed by :class:`EngineCore`. ``state_dict()`` has the reference's keys (``module.``-prefixed
    with more than one worker, like DDP's; ref my_ray_module.py:182,261) and its tensors are views of the flat parameter
    arena."""

    def __init__(self, module: nn.Module, *, device, process_group=None, **core_kw):
        super().__init__()
        if match_reference_mlp(module) is None:
            raise ValueError("EngineModel needs the reference's network shape")
        self.core = EngineCore(module, device=device, process_group=process_group, **core_kw)
        self.module = module.to(device)
        # the wrapped module's parameters become views of the arena: module code (fallback path, user inspection,
        # parameters()) and the engine see the same memory
        with torch.no_grad():
            for n, p in self.module.named_parameters():
                p.data = self.core.view(self.core.p_arena, n)
                p.grad = None
                _OWNER_OF_PARAM[id(p)] = self
        self._optimizer = None
        self._pending: Optional[LazyLoss] = None
        self._eval_seq = 0
        self.world_size = self.core.world

    # ------------------------------------------------------------------------------------------ forward
    def _engine_ok(self, x) -> bool:
        if not isinstance(x, torch.Tensor) or not x.is_cuda:
            return False
        norm = getattr(x, "norm", None) if type(x) is not torch.Tensor else None
        if isinstance(norm, tuple) and norm != (float(self.core.mean), float(self.core.std)):
            return False                 # bytes normalised differently from what the engine folds into layer 1
        x = _raw_u8(x)
        if x.dtype not in (torch.uint8, torch.float32) or x.dim() < 2:
            return False
        return x[0].numel() == self.core.d_in

    def forward(self, x):
        if not self._engine_ok(x):
            return self.module(x)
        if self._pending is not None and self._pending.state != "stepped":
            raise NotImplementedError("a second forward before optimizer.step() (gradient accumulation, several losses) is "
                                      "outside the fused step; use prepare_model(model, fuse='modules')")
        return LazyLogits(self, x, self.training and torch.is_grad_enabled())

    def _loss(self, logits: LazyLogits, y: torch.Tensor):
        """Called by nn.CrossEntropyLoss."""
        plan = self.core.plan(logits.batch)
        x = _raw_u8(logits.x)
        if logits.training:
            plan.discard_staged()
            plan.stage(x, y, after=getattr(logits.x, "_ready", None))
            loss = LazyLoss(self, logits, plan)
            logits._loss = loss
            self._pending = loss
            return loss
        plan.eval_batch(x, y)
        self._eval_seq += 1
        logits._eval_seq = self._eval_seq
        return plan.loss.clone()

    def _step(self) -> bool:
        """Called by optim.SGD.step(): run the fused step when a loss is pending and wanted."""
        loss = self._pending
        if loss is None or loss.state == "stepped":
            return False
        self._pending = None
        if loss.state != "wanted":                 # step() without backward(): torch would update nothing
            loss.plan.discard_staged()
            return True
        loss.step_index = self.core.log_index
        loss.plan.step()
        loss.state = "stepped"
        if self.core.log_index == 3:                 # past the eager warm-up: the launch count of a step is known
            from ..data.loader import TIMING
            k = loss.plan.kernel_launches_per_step
            TIMING["extra"].update(launches_per_train_step=k.get("train_u8", k.get("train_f32", 0)),
                                   gradient_path=f"{self.core.algo}/{self.core.ar_mode}" if self.core.world > 1 else "single GPU")
        return True

    # ------------------------------------------------------------------------------------------ evaluation helpers
    def train(self, mode: bool = True):
        if not mode and self.training:
            self.reset_eval()            # a validation pass starts with model.eval() (ref my_ray_module.py:162)
        return super().train(mode)

    def reset_eval(self):
        self.core.eval_loss_sum.zero_()
        self.core.eval_correct.zero_()

    def read_eval(self):
        return float(self.core.eval_loss_sum.item()), int(self.core.eval_correct.item())

    # ------------------------------------------------------------------------------------------ state
    def state_dict(self, *args, destination=None, prefix: str = "", keep_vars: bool = False):
        sd = self.core.model_state_dict()
        out = destination if destination is not None else {}
        for k, v in sd.items():
            out[prefix + k] = v if keep_vars else v.detach()
        return out

    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):
        names = set(self.core.param_names)
        given = {(k[len("module."):] if k.startswith("module.") else k): v for k, v in state_dict.items()}
        if strict and set(given) != names:
            raise RuntimeError(f"state_dict keys do not match: missing {sorted(names - set(given))}, unexpected {sorted(set(given) - names)}")
        self.core.load_model_state_dict({k: v for k, v in given.items() if k in names})
        return torch.nn.modules.module._IncompatibleKeys([], [])

    def zero_grad(self, set_to_none: bool = True):
        for p in self.module.parameters():
            p.grad = None

    def close(self):
        self.core.close()


class EngineSGD:
    """``optim.SGD`` for an :class:`EngineModel`: ``step()`` fires the fused step; ``state_dict()`` has torch's layout
    (ref my_ray_module.py:142,160,183)."""

    def __init__(self, model: EngineModel, lr: float = 1e-3, momentum: float = 0.0, dampening: float = 0.0,
                 weight_decay: float = 0.0, nesterov: bool = False):
        self.model = model
        core = model.core
        core.lr, core.momentum, core.weight_decay = lr, momentum, weight_decay
        core.sgd.update(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=int(nesterov))
        self.dampening, self.nesterov = dampening, nesterov
        self.param_groups = [dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov,
                                  maximize=False, foreach=None, differentiable=False, fused=None,
                                  params=list(range(len(core.param_names))))]
        self.defaults = dict(lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov)
        model._optimizer = self

    def zero_grad(self, set_to_none: bool = True):
        """Nothing to clear: the fused backward overwrites the gradient arena."""
        self.model.zero_grad()

    def step(self):
        core = self.model.core
        g = self.param_groups[0]
        if g["lr"] != core.lr:                        # LR schedulers write param_groups[0]['lr']
            core.set_lr(g["lr"])
        if self.model._step():
            return
        self._module_path_step()

    @torch.no_grad()
    def _module_path_step(self):
        """Fallback for batches that went through the wrapped module under autograd (inputs the engine does not take):
        gradients are gathered into the arena, reduced, and the same optimizer kernels run."""
        from .. import ops
        core = self.model.core
        named = dict(self.model.module.named_parameters())
        if all(p.grad is None for p in named.values()):
            return
        cur = torch.cuda.current_stream(core.device)
        for n, p in named.items():
            gv = core.view(core.g_arena, n)
            gv.copy_(p.grad) if p.grad is not None else gv.zero_()
        if core.world > 1 and core.ar_mode == "fused":
            core.symm.sgd_step(core.g_off, core.p_off, core.m_off, core.total, core.sgd, counter=core.step_counter,
                               algo=core.algo, stream=cur)
        else:
            if core.world > 1:
                core.symm.all_reduce(core.g_off, core.total, scale=1.0 / core.world, algo=core.algo, stream=cur)
            ops.sgd_step(core.p_arena, core.g_arena, core.m_arena, core.sgd, stream=cur, advance=True, counter=core.step_counter)
        core.steps_done += 1
        core.log_index += 1

    def state_dict(self) -> Dict[str, Any]:
        sd = self.model.core.optimizer_state_dict()
        sd["param_groups"][0].update(dampening=self.dampening, nesterov=self.nesterov)
        return sd

    def load_state_dict(self, sd: Dict[str, Any]):
        self.model.core.load_optimizer_state_dict(sd)
        g = sd["param_groups"][0]
        for k in ("lr", "momentum", "dampening", "weight_decay", "nesterov"):
            if k in g:
                self.param_groups[0][k] = g[k]
        self.dampening, self.nesterov = self.param_groups[0]["dampening"], self.param_groups[0]["nesterov"]
        self.model.core.sgd.update(dampening=self.dampening, nesterov=int(self.nesterov))


class EvalAccumulator:
    """Validation loss / accuracy with the sums kept on the device: ``update(loss, pred, y)`` per batch, ``compute()`` once
    per epoch does the only device->host read (the reference syncs twice per batch, my_ray_module.py:168,170). Works for
    engine-backed models (the eval graph accumulates as a side effect), plain CUDA tensors (fused loss kernel) and CPU."""

    def __init__(self, model: Optional[nn.Module] = None):
        self.batches = 0
        self.samples = 0
        self._owner: Optional[EngineModel] = model if isinstance(model, EngineModel) else None
        if self._owner is not None:
            self._owner.reset_eval()
        self._loss_sum = None
        self._correct = None

    def update(self, loss, pred, y):
        if isinstance(pred, LazyLogits) and not pred.training and pred._eval_seq is not None:
            if self._owner is None:                      # model.eval() zeroed the device sums before this pass
                self._owner = pred.owner
        else:
            pred_t = pred.materialize() if isinstance(pred, LazyLogits) else pred
            if self._loss_sum is None:
                self._loss_sum = torch.zeros((), device=pred_t.device)
                self._correct = torch.zeros((), dtype=torch.int64, device=pred_t.device)
            self._loss_sum += loss.detach() if isinstance(loss, torch.Tensor) else float(loss)
            self._correct += (pred_t.argmax(1) == y).sum()
        self.batches += 1
        self.samples += int(y.shape[0])

    def compute(self):
        """(mean of the per-batch mean losses, accuracy) — the reference's ``test_loss /= len(loader)``, ``correct / total``."""
        loss_sum, correct = 0.0, 0
        if self._owner is not None:
            ls, c = self._owner.read_eval()
            loss_sum += ls
            correct += c
        if self._loss_sum is not None:
            loss_sum += float(self._loss_sum.item())
            correct += int(self._correct.item())
        return loss_sum / max(1, self.batches), correct / max(1, self.samples)
