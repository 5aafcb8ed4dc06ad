"""Whole-step engine for the reference MLP: forward, loss, backward, the NVLink gradient reduction and the
optimizer as ONE replayable CUDA graph of hand-written kernels.

The reference's hot loop (``/root/reference/my_ray_module.py:153-160``) launches ~30 library kernels per
step and two NCCL all-reduces; here a step is (see ``MLPTrainEngine._enqueue_train_step``):

    [staging stream: uint8 -> fp32]  fwd1 · fwd2 · fused head (fwd3 + loss + layer-3 backward + db2) ·
    wgrad2 · dgrad2 (+ column sums) · db1 · wgrad1 ·
    world 1 : fused SGD over the flat arena
    world N : ONE kernel that reduce-scatters the gradient arena through the switch, applies SGD to this rank's
              1/N slice and multicasts the updated parameters to every replica (``symm.sgd_step``)

Two objects:

* :class:`EngineCore` owns everything that is independent of the batch size: the flat parameter / gradient /
  momentum arenas (peer-mapped symmetric memory when distributed, laid out like torch DDP's buckets), the
  optimizer hyper-parameters and counters in device memory, the streams, and the state in the reference's
  checkpoint format.
* :class:`MLPTrainEngine` is the plan for ONE per-GPU batch size on a core: static activation buffers, the
  two input slots and the captured graphs. ``train.torch.prepare_model`` keeps one core per model and one plan
  per batch size it meets (the full batch, the epoch's tail batch, the validation batch).
"""
from __future__ import annotations

import os
from typing import Any, Dict, List, Optional

import torch
import torch.distributed as dist
from torch import nn

from .. import ops
from ..models.mlp import NeuralNetwork
from . import buckets as _buckets

_LOGIT_LD = 16          # logits / dlogits rows padded to 16 floats (TMA needs 16-byte row pitch)
LOSS_LOG_SLOTS = 4096   # ring of per-step losses in pinned host memory (power of two)


def _tf32_exact_pixels(shift: float) -> bool:
    """True when pixel + shift is exactly representable in tf32 (10 explicit mantissa bits) for every uint8 pixel."""
    import numpy as np
    v = np.arange(256, dtype=np.float64) + float(shift)
    f = v.astype(np.float32)
    return bool((f.astype(np.float64) == v).all() and ((f.view(np.uint32) & 0x1FFF) == 0).all())


def match_reference_mlp(model: nn.Module) -> Optional[Dict[str, Any]]:
    """Recognise the reference's network shape (``my_ray_module.py:94-112``): an optional ``Flatten`` plus ONE
    ``nn.Sequential`` of ``Linear, ReLU[, Dropout], Linear, ReLU[, Dropout], Linear, ReLU`` whose sizes the fused
    kernels support. Returns ``{"prefix", "names", "p_drop"}`` (parameter-name prefix of the Sequential, names of
    its three Linear slots) or ``None``. Chains already rewritten by ``fuse_linear_chains`` are recognised too."""
    seqs = [(n, m) for n, m in model.named_children() if isinstance(m, nn.Sequential)]
    others = [m for _, m in model.named_children() if not isinstance(m, (nn.Sequential, nn.Flatten))]
    if isinstance(model, nn.Sequential):
        seqs, others = [("", model)], []
    if len(seqs) != 1 or others:
        return None
    prefix, seq = seqs[0]
    layers, p_drops = [], []
    items = list(seq._modules.items())
    i = 0
    while i < len(items):
        name, m = items[i]
        is_lin = type(m) is nn.Linear or (hasattr(m, "weight") and hasattr(m, "in_features") and hasattr(m, "p_drop"))
        if not is_lin:
            if isinstance(m, nn.Identity):
                i += 1
                continue
            return None
        if m.bias is None or m.weight.dtype != torch.float32:
            return None
        relu = bool(getattr(m, "relu", False))
        p = float(getattr(m, "p_drop", 0.0))
        j = i + 1
        while j < len(items) and isinstance(items[j][1], (nn.ReLU, nn.Dropout, nn.Identity)):
            if isinstance(items[j][1], nn.ReLU):
                relu = True
            elif isinstance(items[j][1], nn.Dropout):
                p = float(items[j][1].p)
            j += 1
        if not relu:
            return None
        layers.append((name, m))
        p_drops.append(p)
        i = j
    if len(layers) != 3 or p_drops[2] != 0.0 or p_drops[0] != p_drops[1]:
        return None
    (_, l1), (_, l2), (_, l3) = layers
    d_in, d_h, d_out = l1.weight.shape[1], l1.weight.shape[0], l3.weight.shape[0]
    if tuple(l2.weight.shape) != (d_h, d_h) or l3.weight.shape[1] != d_h:
        return None
    if d_in % 4 or d_h % 4 or d_h > 512 or d_out not in (10, 16):
        return None
    pre = f"{prefix}." if prefix else ""
    return {"prefix": pre, "names": [f"{pre}{n}" for n, _ in layers], "p_drop": p_drops[0]}


class EngineCore:
    """Batch-size independent state of the fused MLP trainer (see the module docstring)."""

    def __init__(self, model: Optional[nn.Module] = None, *, lr: float = 1e-3, momentum: float = 0.9, weight_decay: float = 0.0,
                 device=None, seed: int = 0x5EED, precision: int = ops.PRECISION_FP32, process_group=None,
                 distributed: Optional[bool] = None, algo: str = "auto", input_mean: float = 0.5, input_std: float = 0.5,
                 fold_input: bool = True, ar_mode: Optional[str] = None, ar_overlap_blocks: Optional[int] = None,
                 fused_head: bool = True):
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        if self.device.type == "cuda" and self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        if self.device.type != "cuda":
            raise ValueError("the fused engine runs the sm_100a kernels and needs a CUDA device; CPU workers use the "
                             "module path (prepare_model + optim.SGD)")
        torch.cuda.set_device(self.device)
        model = model if model is not None else NeuralNetwork()
        spec = match_reference_mlp(model)
        if spec is None:
            raise ValueError("the fused engine expects the reference's network shape: Flatten + Sequential(Linear, ReLU, "
                             "Dropout, Linear, ReLU, Dropout, Linear, ReLU) with hidden <= 512 and 10 or 16 classes")
        self.lin_names: List[str] = spec["names"]
        self.p_drop = float(spec["p_drop"])
        by_name = dict(model.named_parameters())
        w = [by_name[f"{n}.weight"] for n in self.lin_names]
        self.d_in, self.d_h, self.d_out = w[0].shape[1], w[0].shape[0], w[2].shape[0]
        self.seed, self.precision = seed, precision
        self.mean, self.std = input_mean, input_std
        # uint8 batches: the normalisation x = s*pixel + t = s*(pixel + t/s) is folded into layer 1. Raw (or shifted) pixels
        # are exact in tf32, so the two K = 784 GEMMs drop one of their three 3xTF32 passes (include/rtdc_b200.h,
        # rtdc_gemm_desc.a_exact). "shift": pixel + t/s is itself exact (mean 0.5 -> pixel - 127.5), the layer just scales its
        # accumulator by s. "fold": otherwise W x + b = s*(W pixel) + (b + t*rowsum(W)), dW = s*(dY^T pixel) + t*db 1^T.
        self.in_scale, self.in_shift = 1.0 / (255.0 * input_std), -input_mean / input_std
        self.input_mode = "plain"
        if fold_input and precision == ops.PRECISION_FP32:
            self.input_mode = "shift" if _tf32_exact_pixels(self.in_shift / self.in_scale) else "fold"
        self.fold_input = self.input_mode != "plain"
        self.world = dist.get_world_size(process_group) if (dist.is_initialized() and distributed is not False) else 1
        self.rank = dist.get_rank(process_group) if self.world > 1 else 0
        self.group = process_group
        self.algo = algo
        # How the gradients meet the optimizer when distributed:
        #   "fused"   : one kernel = reduce-scatter (NVLS multimem.ld_reduce or P2P loads) + SGD on this rank's 1/world slice +
        #               parameter multicast (multimem.st or P2P stores). No gradient all-gather, no full-arena optimizer pass.
        #   "merged"  : one all-reduce of the whole gradient arena after backward, then the full-arena SGD kernel on every rank.
        #   "bucketed": each DDP bucket is all-reduced on the comm stream as soon as its last gradient exists.
        self.ar_mode = ar_mode or os.environ.get("RTDC_AR_MODE", "fused")
        if self.ar_mode not in ("fused", "merged", "bucketed"):
            raise ValueError(f"ar_mode must be 'fused', 'merged' or 'bucketed' (got {self.ar_mode!r})")
        self.ar_overlap_blocks = int(os.environ.get("RTDC_AR_OVERLAP_BLOCKS", "0")) if ar_overlap_blocks is None else ar_overlap_blocks
        self.side_small = os.environ.get("RTDC_SIDE_SMALL", "1") != "0"
        self.stage_blocks = int(os.environ.get("RTDC_STAGE_BLOCKS", "0"))   # optional grid cap of an overlapped conversion (0 = full speed: measured best)
        self.fused_head = fused_head and self.d_h <= 512 and self.d_h % 4 == 0 and self.d_out in (10, 16)

        # ---- arenas (bucket order == torch DDP's for this model)
        named = [(n, p) for n, p in model.named_parameters()]
        self.buckets, total = _buckets.plan(named)
        self.total = total
        dev = self.device
        self.symm = None
        if self.world > 1:
            from .symm import SymmetricArena
            self.symm = SymmetricArena(3 * (total * 4 + 1024) + 4096, device=dev, group=process_group)
            self.g_off, self.g_arena = self.symm.alloc(total, torch.float32)
            self.p_off, self.p_arena = self.symm.alloc(total, torch.float32)
            self.m_off, self.m_arena = self.symm.alloc(total, torch.float32)
            self.p_arena.zero_(); self.m_arena.zero_(); self.g_arena.zero_()
            self.comm_stream = torch.cuda.Stream(dev)
        else:
            self.g_off = self.p_off = self.m_off = 0
            self.p_arena = torch.zeros(total, device=dev)
            self.m_arena = torch.zeros(total, device=dev)
            self.g_arena = torch.zeros(total, device=dev)
            self.comm_stream = None
        self.slots: Dict[str, _buckets.ParamSlot] = {s.name: s for b in self.buckets for s in b.slots}
        self.param_names = [n for n, _ in named]            # registration order (state_dict / optimizer order)
        by_name = dict(named)
        with torch.no_grad():
            for name in self.slots:
                self.view(self.p_arena, name).copy_(by_name[name].detach().to(dev, torch.float32))
        if self.world > 1:
            self._broadcast_params()
        self.W = [self.view(self.p_arena, f"{n}.weight") for n in self.lin_names]
        self.b = [self.view(self.p_arena, f"{n}.bias") for n in self.lin_names]
        self.dW = [self.view(self.g_arena, f"{n}.weight") for n in self.lin_names]
        self.db = [self.view(self.g_arena, f"{n}.bias") for n in self.lin_names]
        self.bucket_of_layer = [self.slots[f"{n}.weight"].bucket for n in self.lin_names]

        self.b1_eff = torch.zeros(self.d_h, device=dev)      # b1 + t*rowsum(W1) for the folded input layer
        self.loss = torch.zeros((), device=dev)
        self.step_counter = torch.zeros(1, dtype=torch.int32, device=dev)
        self.eval_loss_sum = torch.zeros((), device=dev)
        self.eval_correct = torch.zeros((), dtype=torch.int64, device=dev)
        # every training step's loss also lands in this pinned ring (slot = step counter mod size), written by the optimizer
        # kernel straight over PCIe: the host can follow the loss curve without ever synchronising with the step stream
        self.loss_log = torch.full((LOSS_LOG_SLOTS,), float("nan"), dtype=torch.float32).pin_memory()
        self.sgd = ops.SgdState(dev, lr, momentum, 0.0, weight_decay, False)
        self.lr, self.momentum, self.weight_decay = lr, momentum, weight_decay
        self.steps_done = 0
        self.log_index = 0                  # host mirror of the device step counter (slot of the next loss in loss_log)
        self._momentum_gathered_at = 0      # steps_done when this rank's momentum arena last held every slice
        self.stream = torch.cuda.Stream(dev)
        self.stage_stream = torch.cuda.Stream(dev)
        # small reductions (head partials, db1) run on a side stream next to the wgrad GEMMs, whose 144 CTAs leave four SMs idle
        self.aux_stream = torch.cuda.Stream(dev)
        self.plans: Dict[int, "MLPTrainEngine"] = {}

    # ------------------------------------------------------------------------------------------ helpers
    def view(self, arena: torch.Tensor, name: str) -> torch.Tensor:
        s = self.slots[name]
        return arena[s.offset:s.offset + s.numel].view(s.shape)

    def _broadcast_params(self):
        """C2: every replica starts from rank 0's parameters, pulled over NVLink out of rank 0's parameter arena."""
        torch.cuda.current_stream(self.device).synchronize()
        self.symm.broadcast(self.p_off, self.total * 4, 0, self.p_arena)
        torch.cuda.current_stream(self.device).synchronize()

    def plan(self, batch_size: int, **kw) -> "MLPTrainEngine":
        """The (cached) execution plan for one per-GPU batch size."""
        b = int(batch_size)
        if b not in self.plans:
            self.plans[b] = MLPTrainEngine(core=self, batch_size=b, **kw)
        return self.plans[b]

    def set_lr(self, lr: float):
        self.lr = lr
        self.sgd.update(lr=lr)

    def recent_losses(self, n: int) -> List[float]:
        """The last ``n`` training losses from the pinned ring (synchronises with the step stream first)."""
        self.stream.synchronize()
        n = min(n, self.log_index, LOSS_LOG_SLOTS)
        c = self.log_index
        return [float(self.loss_log[(c - n + i) % LOSS_LOG_SLOTS]) for i in range(n)]

    # ---- state in the reference's checkpoint format ------------------------------------------------
    def model_state_dict(self, prefix: Optional[str] = None) -> Dict[str, torch.Tensor]:
        """Views into the parameter arena, keyed like ``NeuralNetwork.state_dict()``; with more than one
        rank the keys carry DDP's ``module.`` prefix (ref my_ray_module.py:182,261)."""
        pre = ("module." if self.world > 1 else "") if prefix is None else prefix
        return {pre + n: self.view(self.p_arena, n) for n in self.param_names}

    def _gather_momentum(self):
        """With the fused distributed step each rank only advances the momentum of the slice it owns. Pull the other
        slices out of their owners' arenas over NVLink (no collective: any single rank may do this, e.g. rank 0 writing a
        checkpoint; the owners cannot run ahead because the next step's kernel waits for this rank at its first barrier)."""
        if self.world == 1 or self.ar_mode != "fused" or self._momentum_gathered_at == self.steps_done:
            return
        cur = torch.cuda.current_stream(self.device)
        self.stream.wait_stream(cur)
        self.symm.gather_slices(self.m_off, self.total, stream=self.stream)
        cur.wait_stream(self.stream)
        self._momentum_gathered_at = self.steps_done

    def optimizer_state_dict(self) -> Dict[str, Any]:
        state = {}
        if self.momentum != 0 and self.steps_done > 0:
            self._gather_momentum()
            state = {i: {"momentum_buffer": self.view(self.m_arena, n)} for i, n in enumerate(self.param_names)}
        return {"state": state,
                "param_groups": [dict(lr=self.lr, momentum=self.momentum, dampening=0, weight_decay=self.weight_decay,
                                      nesterov=False, maximize=False, foreach=None, differentiable=False, fused=None,
                                      params=list(range(len(self.param_names))))]}

    def load_model_state_dict(self, sd: Dict[str, torch.Tensor]):
        with torch.no_grad():
            for k, v in sd.items():
                name = k[len("module."):] if k.startswith("module.") else k
                self.view(self.p_arena, name).copy_(v)

    def load_optimizer_state_dict(self, sd: Dict[str, Any]):
        g = sd["param_groups"][0]
        self.lr, self.momentum, self.weight_decay = g["lr"], g["momentum"], g["weight_decay"]
        with torch.no_grad():
            for i, st in sd["state"].items():
                self.view(self.m_arena, self.param_names[int(i)]).copy_(st["momentum_buffer"])
        resumed = bool(sd["state"])
        self.sgd.update(lr=self.lr, momentum=self.momentum, weight_decay=self.weight_decay, first_step=0 if resumed else 1)
        if resumed:
            self.steps_done = max(self.steps_done, 1)
            self._momentum_gathered_at = self.steps_done

    def copy_to_module(self, model: nn.Module):
        model.load_state_dict({k: v.detach().clone() for k, v in self.model_state_dict(prefix="").items()})
        return model

    def close(self):
        for p in self.plans.values():
            p._graphs.clear()
        self.plans.clear()
        if self.symm is not None:
            torch.cuda.synchronize(self.device)
            self.symm.close()
            self.symm = None


_CORE_KW = ("lr", "momentum", "weight_decay", "device", "seed", "precision", "process_group", "distributed", "algo", "input_mean",
            "input_std", "fold_input", "ar_mode", "ar_overlap_blocks", "fused_head")


class MLPTrainEngine:
    """Execution plan of the fused training / evaluation step for one per-GPU batch size.

    ``MLPTrainEngine(model, batch_size=B, ...)`` builds a private :class:`EngineCore`; ``core.plan(B)`` adds a plan to an
    existing core (same parameters, optimizer state, counters and streams)."""

    def __init__(self, model: Optional[nn.Module] = None, *, batch_size: int, use_graph: bool = True,
                 core: Optional[EngineCore] = None, **core_kw):
        unknown = set(core_kw) - set(_CORE_KW)
        if unknown:
            raise TypeError(f"unexpected arguments {sorted(unknown)}")
        self._owns_core = core is None
        if core is None:
            core = EngineCore(model, **core_kw)
            core.plans[int(batch_size)] = self
        self.core = c = core
        # the plan reads the core's state through these aliases (same tensors)
        for k in ("device", "p_drop", "d_in", "d_h", "d_out", "seed", "precision", "mean", "std", "in_scale", "in_shift", "input_mode",
                  "fold_input", "world", "rank", "group", "algo", "ar_mode", "ar_overlap_blocks", "side_small", "stage_blocks",
                  "fused_head", "buckets", "total", "symm", "g_off", "p_off", "m_off", "p_arena", "m_arena", "g_arena", "comm_stream",
                  "slots", "param_names", "W", "b", "dW", "db", "bucket_of_layer", "b1_eff", "loss", "step_counter", "eval_loss_sum",
                  "eval_correct", "loss_log", "sgd", "stream", "stage_stream", "aux_stream"):
            setattr(self, k, getattr(c, k))
        self.B = int(batch_size)
        dev = self.device

        # ---- static activations (graph replays need fixed addresses)
        B = self.B
        # two input slots: batch i+1 is staged (copied in and converted) on its own stream while step i runs
        self.x_u8 = [torch.zeros(B, self.d_in, dtype=torch.uint8, device=dev) for _ in range(2)]
        self.x = [torch.zeros(B, self.d_in, device=dev) for _ in range(2)]
        self.y = [torch.zeros(B, dtype=torch.int64, device=dev) for _ in range(2)]
        self._ready: List[Optional[torch.cuda.Event]] = [None, None]   # slot staged
        self._free: List[Optional[torch.cuda.Event]] = [None, None]    # last consumer of the slot done
        self._staged: List[tuple] = []                                  # FIFO of (slot, from_u8)
        self._next_slot = 0
        self._stage_done: Optional[torch.cuda.Event] = None
        self.h1 = torch.empty(B, self.d_h, device=dev)
        self.h2 = torch.empty(B, self.d_h, device=dev)
        self.z = torch.zeros(B, _LOGIT_LD, device=dev)
        self.dz = torch.zeros(B, _LOGIT_LD, device=dev)
        self.dh2 = torch.empty(B, self.d_h, device=dev)
        self.dh1 = torch.empty(B, self.d_h, device=dev)
        self.dh1_colsum = torch.zeros(4 * ((B + 127) // 128), self.d_h, device=dev)
        self.h1_bits = torch.zeros(B, (self.d_h + 31) // 32, dtype=torch.int32, device=dev)   # (h1 > 0), 1 bit per element
        self.use_graph = use_graph
        self._graphs: Dict[tuple, torch.cuda.CUDAGraph] = {}
        self._warm: Dict[str, int] = {}
        self.kernel_launches_per_step = {}

    # the optimizer's scalars live on the core (all plans of a model share them)
    steps_done = property(lambda self: self.core.steps_done, lambda self, v: setattr(self.core, "steps_done", v))
    lr = property(lambda self: self.core.lr)
    momentum = property(lambda self: self.core.momentum)
    weight_decay = property(lambda self: self.core.weight_decay)

    # ------------------------------------------------------------------------------------------ helpers
    def _view(self, arena: torch.Tensor, name: str) -> torch.Tensor:
        return self.core.view(arena, name)

    def _reduce_bucket(self, b: int, main: torch.cuda.Stream, last: bool = False):
        if self.world == 1 or self.ar_mode != "bucketed":
            return
        bk = self.buckets[b]
        ev = torch.cuda.Event()
        ev.record(main)
        self.comm_stream.wait_event(ev)
        self.symm.all_reduce(self.g_off + bk.offset * 4, bk.numel, scale=1.0 / self.world, algo=self.algo,
                             max_blocks=0 if last else self.ar_overlap_blocks, stream=self.comm_stream)

    # ------------------------------------------------------------------------------------------ the step
    def _convert_input(self, src_u8: torch.Tensor, slot: int, s, paced: bool) -> int:
        """uint8 pixels -> the fp32 operand layer 1 consumes (``self.x[slot]``), on stream ``s``. ``paced``: the
        conversion overlaps a running step, so its grid is capped (``stage_blocks``) to stay out of the GEMMs' way."""
        blocks = self.stage_blocks if paced else 0
        if self.input_mode == "shift":
            ops.u8_affine(src_u8, shift=self.in_shift / self.in_scale, out=self.x[slot], blocks=blocks, stream=s)   # pixel - 127.5
        elif self.input_mode == "fold":
            ops.u8_affine(src_u8, out=self.x[slot], blocks=blocks, stream=s)                                        # raw pixels
        else:
            ops.u8_affine(src_u8, scale=self.in_scale, shift=self.in_shift, out=self.x[slot], blocks=blocks, stream=s)
        return 1

    def _forward(self, s, train: bool, from_u8: bool, slot: int, with_logits: bool = True):
        n = 0
        p = self.p_drop if train else 0.0
        kw = dict(seed=self.seed, step_tensor=self.step_counter, precision=self.precision, stream=s)
        x = self.x[slot]
        kw1 = dict(kw, mask_bits_out=self.h1_bits if train else None)   # dgrad2 masks with 1 bit per element of h1
        if from_u8 and self.input_mode == "shift":
            ops.linear_forward(x, self.W[0], self.b[0], relu=True, drop_p=p, salt=1, out=self.h1, x_exact=True,
                               alpha=self.in_scale, **kw1); n += 1
        elif from_u8 and self.input_mode == "fold":
            ops.fold_input_bias(self.W[0], self.b[0], self.in_shift, out=self.b1_eff, stream=s); n += 1
            ops.linear_forward(x, self.W[0], self.b1_eff, relu=True, drop_p=p, salt=1, out=self.h1, x_exact=True,
                               alpha=self.in_scale, **kw1); n += 1
        else:
            ops.linear_forward(x, self.W[0], self.b[0], relu=True, drop_p=p, salt=1, out=self.h1, **kw1); n += 1
        ops.linear_forward(self.h1, self.W[1], self.b[1], relu=True, drop_p=p, salt=2, out=self.h2, **kw); n += 1
        if with_logits:
            ops.linear_forward(self.h2, self.W[2], self.b[2], relu=True, out=self.z, n_store=_LOGIT_LD, **kw); n += 1
        return n

    def _enqueue_train_step(self, s: torch.cuda.Stream, from_u8: bool, slot: int) -> int:
        B, prec = self.B, self.precision
        inv_keep = 1.0 / (1.0 - self.p_drop) if self.p_drop > 0 else 1.0
        fused_head = self.fused_head
        side = self.side_small and (self.world == 1 or self.ar_mode != "bucketed")
        aux = self.aux_stream
        n = self._forward(s, True, from_u8, slot, with_logits=not fused_head)
        x, y = self.x[slot], self.y[slot]
        if fused_head:
            # layer 3 forward + loss + layer 3 backward in one pass over h2 (exact fp32 on the CUDA cores); also emits db2
            hk = dict(dh=self.dh2, dw=self.dW[2], db=self.db[2], loss=self.loss, grad_scale=1.0 / B, mask_scale=inv_keep,
                      db_below=self.db[1])
            if side:
                ops.mlp_head_train(self.h2, self.W[2], self.b[2], y, stream=s, part=1, **hk)
                aux.wait_stream(s)                                   # the partials' reduction overlaps wgrad2
                ops.mlp_head_train(self.h2, self.W[2], self.b[2], y, stream=aux, part=2, scratch_stream=s, **hk)
            else:
                ops.mlp_head_train(self.h2, self.W[2], self.b[2], y, stream=s, **hk)
            n += 2
        else:
            ops.softmax_cross_entropy(self.z, y, classes=self.d_out, dlogits=self.dz, grad_scale=1.0 / B, relu_mask=True,
                                      loss_out=self.loss, dbias_out=self.db[2], stream=s); n += 1   # also emits db3
            ops.linear_wgrad(self.dz, self.h2, N_out=self.d_out, out=self.dW[2], precision=prec, stream=s); n += 2
            ops.linear_dgrad(self.dz, self.W[2], N_out=self.d_out, mask_src=self.h2, mask_scale=inv_keep, out=self.dh2,
                             precision=prec, stream=s); n += 1
        # layer 2
        ops.linear_wgrad(self.dh2, self.h1, out=self.dW[1], precision=prec, stream=s); n += 2
        if not fused_head:
            ops.colsum(self.dh2, out=self.db[1], stream=s); n += 1
        launched = set()
        if self.bucket_of_layer[1] != self.bucket_of_layer[0]:
            self._reduce_bucket(self.bucket_of_layer[1], s); launched.add(self.bucket_of_layer[1])
            n += self.world > 1 and self.ar_mode == "bucketed"
        # dgrad2 also leaves per-32-row column sums of dh1 (592 x 512 instead of 18944 x 512 to sum for db1)
        ops.linear_dgrad(self.dh2, self.W[1], mask_bits=self.h1_bits, mask_scale=inv_keep, out=self.dh1, precision=prec,
                         colsum_partial=self.dh1_colsum, stream=s); n += 1
        # layer 1 (its input needs no gradient); with folded input dW1 = s*(dh1^T raw) + t*db1 1^T
        side_cs = side and not (from_u8 and self.input_mode == "fold")     # the folded wgrad1 epilogue reads db1
        if side_cs:
            aux.wait_stream(s)
            ops.colsum(self.dh1_colsum, out=self.db[0], stream=aux); n += 1      # overlaps wgrad1
        else:
            ops.colsum(self.dh1_colsum, out=self.db[0], stream=s); n += 1
        if from_u8 and self.input_mode == "shift":
            ops.linear_wgrad(self.dh1, x, out=self.dW[0], precision=prec, x_exact=True, alpha=self.in_scale, stream=s); n += 2
        elif from_u8 and self.input_mode == "fold":
            ops.linear_wgrad(self.dh1, x, out=self.dW[0], precision=prec, x_exact=True, alpha=self.in_scale,
                             row_bias=self.db[0], row_bias_scale=self.in_shift, stream=s); n += 2
        else:
            ops.linear_wgrad(self.dh1, x, out=self.dW[0], precision=prec, stream=s); n += 2
        for b in range(len(self.buckets)):
            if b not in launched:
                self._reduce_bucket(b, s, last=True); n += self.world > 1 and self.ar_mode == "bucketed"
        if side:
            s.wait_stream(aux)
        if self.world > 1 and self.ar_mode == "fused":
            # reduce-scatter + SGD on the owned slice + parameter multicast: one kernel, also advances the counters
            self.symm.sgd_step(self.g_off, self.p_off, self.m_off, self.total, self.sgd, counter=self.step_counter, loss=self.loss,
                               loss_log=self.loss_log, algo=self.algo, stream=s); n += 1
            return n
        if self.world > 1 and self.ar_mode == "merged":
            self.symm.all_reduce(self.g_off, self.total, scale=1.0 / self.world, algo=self.algo, stream=s); n += 1
        elif self.world > 1:
            s.wait_stream(self.comm_stream)
        ops.sgd_step(self.p_arena, self.g_arena, self.m_arena, self.sgd, stream=s, advance=True, counter=self.step_counter,
                     loss=self.loss, loss_log=self.loss_log); n += 1
        return n

    def _enqueue_eval(self, s: torch.cuda.Stream, from_u8: bool, slot: int) -> int:
        n = self._forward(s, False, from_u8, slot)
        ops.softmax_cross_entropy(self.z, self.y[slot], classes=self.d_out, loss_out=self.loss, loss_sum_accum=self.eval_loss_sum,
                                  correct_accum=self.eval_correct, stream=s)
        return n + 1

    def _run(self, key: str, slot: int, enqueue):
        """Two eager warm-up executions (allocations, scratch, first-step flag), then capture (both input slots at
        once, so no capture falls into a later step) + replay."""
        s = self.stream
        cur = torch.cuda.current_stream(self.device)
        s.wait_stream(cur)
        if self._ready[slot] is not None:
            s.wait_event(self._ready[slot])
        with torch.cuda.stream(s):
            if not self.use_graph or self._warm.get(key, 0) < 2:
                self.kernel_launches_per_step[key] = enqueue(s, slot) + (1 if key.endswith("u8") else 0)   # + the staging kernel
                self._warm[key] = self._warm.get(key, 0) + 1
            else:
                if (key, slot) not in self._graphs:
                    if self.world > 1:
                        dist.barrier(group=self.group)      # all ranks capture (and first replay) together
                    for sl in (0, 1):
                        g = torch.cuda.CUDAGraph()
                        with torch.cuda.graph(g, stream=s, capture_error_mode="thread_local"):
                            enqueue(s, sl)
                        self._graphs[(key, sl)] = g
                self._graphs[(key, slot)].replay()
            ev = torch.cuda.Event()
            ev.record(s)
            self._free[slot] = ev
        cur.wait_stream(s)
        if self._stage_done is not None:
            cur.wait_event(self._stage_done)     # orders the caller's stream after the staging reads of its tensors

    # ------------------------------------------------------------------------------------------ public API
    def stage(self, x: torch.Tensor, y: torch.Tensor, _paced: bool = True, after: Optional[torch.cuda.Event] = None) -> None:
        """Copy a batch in (H2D from pinned host memory, or straight from a device tensor) and convert it for layer 1
        on the staging stream. Call it for batch i+1 *before* ``step()`` of batch i: the copy and the uint8 -> fp32
        conversion then run concurrently with that step instead of in front of the next one. At most two batches can
        be staged. ``after``: the event that marks ``x`` / ``y`` as produced (a loader's copy-done event); without it the
        staging stream waits for everything queued on the caller's stream — which, right after a ``step()``, includes that
        step, and the conversion would then run behind it instead of next to it."""
        if x.shape[0] != self.B:
            raise ValueError(f"engine was built for batches of {self.B} samples (got {x.shape[0]})")
        if len(self._staged) >= 2:
            raise RuntimeError("two batches are already staged; call step() first")
        slot = self._next_slot
        self._next_slot ^= 1
        st = self.stage_stream
        if after is not None:
            st.wait_event(after)
        else:
            st.wait_stream(torch.cuda.current_stream(self.device))     # x / y were produced on the caller's stream
        if self._free[slot] is not None:
            st.wait_event(self._free[slot])                          # the step that last read this slot is done
        from_u8 = x.dtype == torch.uint8
        with torch.cuda.stream(st):
            if after is None:
                for t in (x, y):
                    if t.is_cuda:
                        t.record_stream(st)                          # the caller may drop its reference right away
            # (with `after` the batch lives in a loader's ring buffer, which is reused only after the loader has seen the
            #  consumer's stream pass this staging, see U8BatchLoader._free_ev)
            x2 = x.reshape(self.B, -1)
            self.y[slot].copy_(y, non_blocking=True)
            if from_u8:
                if x2.is_cuda and x2.is_contiguous() and x2.data_ptr() % 16 == 0:
                    src = x2                                         # convert straight out of the caller's buffer
                else:
                    self.x_u8[slot].copy_(x2, non_blocking=True)
                    src = self.x_u8[slot]
                self._convert_input(src, slot, st, paced=_paced)
            else:
                self.x[slot].copy_(x2, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(st)
        self._ready[slot] = ev
        self._stage_done = ev
        self._staged.append((slot, from_u8))

    def _consume(self, x, y):
        if x is not None:
            self.stage(x, y, _paced=False)       # nothing to overlap with: convert at full speed
        if not self._staged:
            raise RuntimeError("no staged batch: call stage(x, y) first or pass the batch to step()")
        return self._staged.pop(0)

    def step(self, x: Optional[torch.Tensor] = None, y: Optional[torch.Tensor] = None) -> torch.Tensor:
        """One optimizer step on a batch (uint8 pixels or normalised float32): either the batch passed in, or the
        oldest one handed to :meth:`stage`. Returns the device-resident mean loss of this rank's batch
        (``.item()`` to read it back)."""
        slot, from_u8 = self._consume(x, y)
        self._run("train_u8" if from_u8 else "train_f32", slot, lambda s, sl: self._enqueue_train_step(s, from_u8, sl))
        self.core.steps_done += 1
        self.core.log_index += 1
        return self.loss

    def discard_staged(self):
        """Forget batches that were staged but never consumed by a step."""
        self._staged.clear()

    def forward_loss(self, x: Optional[torch.Tensor] = None, y: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Training-mode forward + loss only (same dropout masks as the step that follows, since the step counter only
        advances in the optimizer kernel): what ``loss.item()`` before ``optimizer.step()`` costs on the lazy API path.
        The batch stays staged for :meth:`step`."""
        if x is not None:
            self.stage(x, y, _paced=False)
        if not self._staged:
            raise RuntimeError("no staged batch")
        slot, from_u8 = self._staged[0]
        s = self.stream
        cur = torch.cuda.current_stream(self.device)
        s.wait_stream(cur)
        s.wait_event(self._ready[slot])
        with torch.cuda.stream(s):
            self._forward(s, True, from_u8, slot)
            ops.softmax_cross_entropy(self.z, self.y[slot], classes=self.d_out, loss_out=self.loss, stream=s)
        cur.wait_stream(s)
        return self.loss

    def eval_batch(self, x: Optional[torch.Tensor] = None, y: Optional[torch.Tensor] = None) -> torch.Tensor:
        """Forward + loss on a batch, accumulating ``eval_loss_sum`` / ``eval_correct`` on the device."""
        slot, from_u8 = self._consume(x, y)
        self._run("eval_u8" if from_u8 else "eval_f32", slot, lambda s, sl: self._enqueue_eval(s, from_u8, sl))
        return self.loss

    def reset_eval(self):
        self.eval_loss_sum.zero_()
        self.eval_correct.zero_()

    def read_eval(self):
        """(sum of per-batch mean losses, correct predictions): one D2H per call (ref: two per batch)."""
        return float(self.eval_loss_sum.item()), int(self.eval_correct.item())

    def predict(self, x: torch.Tensor, train: bool = False) -> torch.Tensor:
        """Logits ``[B, classes]`` for a batch (eval mode unless ``train``)."""
        slot, from_u8 = self._consume(x, torch.zeros(self.B, dtype=torch.int64, device=self.device))
        s = torch.cuda.current_stream(self.device)
        s.wait_event(self._ready[slot])
        self._forward(s, train, from_u8, slot)
        ev = torch.cuda.Event(); ev.record(s); self._free[slot] = ev
        return self.z[:, :self.d_out].clone()

    def set_lr(self, lr: float):
        self.core.set_lr(lr)

    # ---- state (kept on the core) -------------------------------------------------------------------
    def model_state_dict(self, prefix: Optional[str] = None):
        return self.core.model_state_dict(prefix)

    def optimizer_state_dict(self):
        return self.core.optimizer_state_dict()

    def load_model_state_dict(self, sd):
        self.core.load_model_state_dict(sd)

    def load_optimizer_state_dict(self, sd):
        self.core.load_optimizer_state_dict(sd)

    def copy_to_module(self, model: nn.Module):
        return self.core.copy_to_module(model)

    def close(self):
        self._graphs.clear()
        if self._owns_core:
            self.core.close()
