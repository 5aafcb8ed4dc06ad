"""Symmetric memory and NVLink collectives (Python face of ``csrc/symm_mem.cpp`` + ``csrc/allreduce.cu``).

``SymmetricArena`` is one peer-mapped device buffer per rank. Tensors carved out of it
(``arena.tensor(...)``) are ordinary CUDA tensors on this rank and are directly readable/writable by
the all-reduce / broadcast kernels running on the other GPUs. This is what replaces the NCCL process
group that the reference gets implicitly from ``ray.train.torch.prepare_model`` →
``DistributedDataParallel`` (``/root/reference/my_ray_module.py:135``).
"""
from __future__ import annotations

import ctypes as C
import os
import tempfile
import uuid
from typing import Optional

import torch

from ..ops import _lib

ALGO_AUTO, ALGO_ONE_SHOT, ALGO_TWO_SHOT, ALGO_NVLS = 0, 1, 2, 3
_ALGO_NAMES = {"auto": 0, "one_shot": 1, "two_shot": 2, "nvls": 3}
MAX_WORLD = 8


class SymmInfo(C.Structure):
    _fields_ = [("rank", C.c_int), ("world", C.c_int), ("device", C.c_int), ("bytes", C.c_size_t),
                ("local_ptr", C.c_void_p), ("multicast_ptr", C.c_void_p), ("peer_ptrs", C.c_void_p * MAX_WORLD),
                ("peer_table_dev", C.c_void_p), ("pad_table_dev", C.c_void_p), ("epochs_dev", C.c_void_p)]


def _declare(lib):
    if getattr(lib, "_symm_declared", False):
        return lib
    vp, i, sz, f = C.c_void_p, C.c_int, C.c_size_t, C.c_float
    lib.rtdc_symm_create.argtypes = [i, i, i, sz, C.c_char_p, i, C.c_double, C.POINTER(vp)]
    lib.rtdc_symm_bind_multicast.argtypes = [vp]
    lib.rtdc_symm_info.argtypes = [vp, C.POINTER(SymmInfo)]
    lib.rtdc_symm_destroy.argtypes = [vp]
    lib.rtdc_symm_all_reduce.argtypes = [vp, sz, sz, i, f, i, vp, i, vp]
    lib.rtdc_symm_broadcast.argtypes = [vp, sz, sz, i, vp, vp]
    lib.rtdc_symm_barrier.argtypes = [vp, vp]
    lib.rtdc_symm_sgd_step.argtypes = [vp, sz, sz, sz, sz, i, vp, vp, vp, vp, C.c_uint32, i, vp]
    lib.rtdc_symm_gather_slices.argtypes = [vp, sz, sz, vp]
    for n in ("rtdc_symm_create", "rtdc_symm_bind_multicast", "rtdc_symm_info", "rtdc_symm_destroy",
              "rtdc_symm_all_reduce", "rtdc_symm_broadcast", "rtdc_symm_barrier", "rtdc_symm_sgd_step", "rtdc_symm_gather_slices"):
        getattr(lib, n).restype = i
    lib._symm_declared = True
    return lib


class _RawCudaBuffer:
    """Minimal ``__cuda_array_interface__`` carrier so torch can view driver-mapped memory."""

    def __init__(self, ptr: int, nbytes: int, owner):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2}
        self._owner = owner


class SymmetricArena:
    """``nbytes`` of device memory on every rank, each mapped into all ranks' address spaces.

    Collective construction: every rank of the process group must call it with the same size. The
    rendezvous (a unix-socket path prefix) is agreed through ``torch.distributed``.
    """

    def __init__(self, nbytes: int, *, device: Optional[torch.device] = None, group=None, multicast: bool = True,
                 timeout_s: float = 120.0):
        import torch.distributed as dist
        self.lib = _declare(_lib.load())
        self.group = group
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        if self.world > MAX_WORLD:
            raise ValueError(f"SymmetricArena supports up to {MAX_WORLD} ranks on one NVSwitch node (got {self.world})")
        self.device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
        prefix = [None]
        if self.rank == 0:
            prefix[0] = os.path.join(tempfile.gettempdir(), f"rtdc_symm_{os.getpid()}_{uuid.uuid4().hex[:8]}")
        if self.world > 1:
            dist.broadcast_object_list(prefix, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        handle = C.c_void_p()
        _lib.check(self.lib.rtdc_symm_create(self.rank, self.world, self.device.index or 0, int(nbytes),
                                             prefix[0].encode(), int(multicast and self.world > 1), float(timeout_s),
                                             C.byref(handle)))
        self.handle = handle
        if self.world > 1:
            dist.barrier(group=group)                      # every rank has added its device to the multicast object
        _lib.check(self.lib.rtdc_symm_bind_multicast(self.handle))
        info = SymmInfo()
        _lib.check(self.lib.rtdc_symm_info(self.handle, C.byref(info)))
        self.info = info
        self.nbytes = int(info.bytes)
        mc = [1 if info.multicast_ptr else 0]
        if self.world > 1:
            # NVLS is only used when every rank managed to bind
            t = torch.tensor(mc, device=self.device, dtype=torch.int32)
            dist.all_reduce(t, op=dist.ReduceOp.MIN, group=group)
            mc = [int(t.item())]
        self.has_multicast = bool(mc[0])
        self._bytes = torch.as_tensor(_RawCudaBuffer(int(info.local_ptr), self.nbytes, self), device=self.device)
        self._cursor = 0

    # ------------------------------------------------------------------ carving
    def tensor(self, offset: int, numel: int, dtype=torch.float32) -> torch.Tensor:
        nb = numel * torch.empty((), dtype=dtype).element_size()
        if offset % 16 or offset + nb > self.nbytes:
            raise ValueError(f"bad arena range offset={offset} bytes={nb} (arena {self.nbytes})")
        return self._bytes[offset:offset + nb].view(dtype)

    def alloc(self, numel: int, dtype=torch.float32, align: int = 512):
        """Bump-allocate a tensor; returns (offset_bytes, tensor). Same call sequence on all ranks."""
        off = (self._cursor + align - 1) // align * align
        t = self.tensor(off, numel, dtype)
        self._cursor = off + t.numel() * t.element_size()
        return off, t

    # ------------------------------------------------------------------ collectives
    def all_reduce(self, offset: int, numel: int, *, dtype=torch.float32, scale: float = 1.0, algo="auto",
                   out: Optional[torch.Tensor] = None, max_blocks: int = 0, stream=None):
        """In-place sum over ranks of ``numel`` elements at byte ``offset``, times ``scale`` (pass
        ``1/world`` for the DDP mean). ``algo='one_shot'`` writes to ``out`` (a local tensor) instead."""
        a = _ALGO_NAMES[algo] if isinstance(algo, str) else int(algo)
        if a == ALGO_NVLS and not self.has_multicast:
            raise RuntimeError("NVLS all-reduce requested but multicast is unavailable on this system")
        if a == ALGO_AUTO and not self.has_multicast:
            a = ALGO_TWO_SHOT
        dt = {torch.float32: 0, torch.bfloat16: 1}[dtype]
        _lib.check(self.lib.rtdc_symm_all_reduce(self.handle, offset, numel, dt, scale, a,
                                                 out.data_ptr() if out is not None else None, max_blocks,
                                                 _lib.stream_ptr(stream)))

    def sgd_step(self, g_offset: int, p_offset: int, m_offset: int, numel: int, state, *, counter: Optional[torch.Tensor] = None,
                 loss: Optional[torch.Tensor] = None, loss_log: Optional[torch.Tensor] = None, algo="auto", max_blocks: int = 0,
                 stream=None):
        """The distributed optimizer step in one kernel: reduce-scatter of the gradient arena (through the switch with
        NVLS, else peer loads), SGD/momentum on this rank's 1/world slice, updated parameters stored into every replica
        (``rtdc_symm_sgd_step``). ``state`` is an ``ops.SgdState``; offsets are byte offsets of fp32 arenas of ``numel``
        elements inside this arena."""
        a = _ALGO_NAMES[algo] if isinstance(algo, str) else int(algo)
        if a == ALGO_NVLS and not self.has_multicast:
            raise RuntimeError("NVLS requested but multicast is unavailable on this system")
        if a == ALGO_AUTO and not self.has_multicast:
            a = ALGO_TWO_SHOT                       # else the library picks by size (P2P below 64 MiB, see csrc/allreduce.cu)
        if loss_log is not None and not loss_log.is_pinned():
            raise ValueError("loss_log must be pinned host memory")
        _lib.check(self.lib.rtdc_symm_sgd_step(self.handle, g_offset, p_offset, m_offset, numel, a, state.dev.data_ptr(),
                                               counter.data_ptr() if counter is not None else None,
                                               loss.data_ptr() if loss is not None else None,
                                               loss_log.data_ptr() if loss_log is not None else None,
                                               loss_log.numel() if loss_log is not None else 0, max_blocks,
                                               _lib.stream_ptr(stream)))
        state.host.first_step = 0

    def gather_slices(self, offset: int, numel: int, stream=None):
        """Complete this rank's copy of a rank-sliced fp32 arena (e.g. the momentum of :meth:`sgd_step`) by pulling the other
        ranks' slices over NVLink. Not a collective."""
        _lib.check(self.lib.rtdc_symm_gather_slices(self.handle, offset, numel, _lib.stream_ptr(stream)))

    def broadcast(self, offset: int, nbytes: int, root: int, dst: torch.Tensor, stream=None):
        _lib.check(self.lib.rtdc_symm_broadcast(self.handle, offset, nbytes, root, dst.data_ptr(), _lib.stream_ptr(stream)))

    def barrier(self, stream=None):
        _lib.check(self.lib.rtdc_symm_barrier(self.handle, _lib.stream_ptr(stream)))

    def close(self):
        if getattr(self, "handle", None):
            self._bytes = None
            self.lib.rtdc_symm_destroy(self.handle)
            self.handle = None
