"""ctypes binding of ``librtdc_b200.so`` (declarations: ``include/rtdc_b200.h``).

There is no PyTorch fallback for the CUDA ops: on a machine with a GPU a missing or stale library
is an error, not a silent eager path.
"""
from __future__ import annotations

import ctypes as C
import threading
from pathlib import Path

from . import _build

_lock = threading.Lock()
_lib = None


class RtdcError(RuntimeError):
    pass


class GemmDesc(C.Structure):
    _fields_ = [
        ("a", C.c_void_p), ("lda", C.c_int), ("a_mn", C.c_int),
        ("b", C.c_void_p), ("ldb", C.c_int), ("b_mn", C.c_int),
        ("M", C.c_int), ("N", C.c_int), ("K", C.c_int),
        ("out", C.c_void_p), ("ldo", C.c_int), ("n_store", C.c_int),
        ("bias", C.c_void_p), ("relu", C.c_int),
        ("drop_p", C.c_float),
        ("seed", C.c_uint64),
        ("step_ptr", C.c_void_p),
        ("step_imm", C.c_uint32),
        ("salt", C.c_uint32),
        ("mask_src", C.c_void_p), ("ldm", C.c_int), ("mask_scale", C.c_float),
        ("precision", C.c_int),
        ("bn", C.c_int),
        ("splits", C.c_int),
        ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
        ("colsum_partial", C.c_void_p), ("ld_colsum", C.c_int),
        ("a_exact", C.c_int), ("b_exact", C.c_int),
        ("alpha", C.c_float),
        ("row_bias", C.c_void_p), ("row_bias_scale", C.c_float),
        ("mask_bits_out", C.c_void_p), ("ld_mask_out", C.c_int),
        ("mask_bits_in", C.c_void_p), ("ld_mask_in", C.c_int),
    ]


class SgdHyper(C.Structure):
    _fields_ = [("lr", C.c_float), ("momentum", C.c_float), ("dampening", C.c_float),
                ("weight_decay", C.c_float), ("grad_scale", C.c_float), ("nesterov", C.c_int),
                ("first_step", C.c_int), ("pad", C.c_int)]


class AdamHyper(C.Structure):
    _fields_ = [("lr", C.c_float), ("beta1", C.c_float), ("beta2", C.c_float), ("eps", C.c_float),
                ("weight_decay", C.c_float), ("grad_scale", C.c_float), ("decoupled", C.c_int),
                ("step", C.c_int)]


def _declare(lib):
    vp, i, f, sz = C.c_void_p, C.c_int, C.c_float, C.c_size_t
    lib.rtdc_last_error.restype = C.c_char_p
    lib.rtdc_version.restype = i
    lib.rtdc_arch.restype = C.c_char_p
    lib.rtdc_gemm.argtypes = [C.POINTER(GemmDesc), vp]
    lib.rtdc_gemm_scratch_bytes.argtypes = [C.POINTER(GemmDesc)]
    lib.rtdc_gemm_scratch_bytes.restype = sz
    lib.rtdc_gemm_plan.argtypes = [C.POINTER(GemmDesc), C.POINTER(i * 5)]
    lib.rtdc_softmax_ce_scratch_bytes.argtypes = [i]
    lib.rtdc_softmax_ce_scratch_bytes.restype = sz
    lib.rtdc_softmax_ce.argtypes = [vp, i, vp, i, i, vp, vp, f, i, vp, vp, vp, vp, sz, vp]
    lib.rtdc_mlp_head_scratch_bytes.argtypes = [i, i, i]
    lib.rtdc_mlp_head_scratch_bytes.restype = sz
    lib.rtdc_mlp_head_train.argtypes = [vp, i, vp, vp, vp, i, i, i, f, f, vp, i, vp, vp, vp, vp, vp, sz, vp]
    lib.rtdc_mlp_head_train.restype = i
    lib.rtdc_mlp_head_train_part.argtypes = [vp, i, vp, vp, vp, i, i, i, f, f, vp, i, vp, vp, vp, vp, vp, sz, vp, i]
    lib.rtdc_mlp_head_train_part.restype = i
    lib.rtdc_colsum_scratch_bytes.argtypes = [i, i]
    lib.rtdc_colsum_scratch_bytes.restype = sz
    lib.rtdc_colsum.argtypes = [vp, i, i, i, vp, vp, sz, vp]
    lib.rtdc_u8_normalize.argtypes = [vp, vp, sz, f, f, vp]
    lib.rtdc_u8_affine.argtypes = [vp, vp, sz, f, f, vp]
    lib.rtdc_u8_affine_paced.argtypes = [vp, vp, sz, f, f, i, vp]
    lib.rtdc_u8_affine_paced.restype = i
    lib.rtdc_fold_input_bias.argtypes = [vp, i, i, i, vp, f, vp, vp]
    lib.rtdc_u8_gather_rows.argtypes = [vp, vp, i, i, vp, vp, vp, vp]
    lib.rtdc_u8_gather_rows.restype = i
    lib.rtdc_h2d_pieces.argtypes = [vp, vp, vp, vp, i, vp]
    lib.rtdc_h2d_pieces.restype = i
    lib.rtdc_act_backward.argtypes = [vp, i, vp, i, i, i, i, f, vp, i, vp]
    lib.rtdc_act_backward.restype = i
    lib.rtdc_sgd_step.argtypes = [vp, vp, vp, sz, vp, vp]
    lib.rtdc_sgd_step_advance.argtypes = [vp, vp, vp, sz, vp, vp, vp]
    lib.rtdc_sgd_step_advance.restype = i
    lib.rtdc_sgd_step_advance_log.argtypes = [vp, vp, vp, sz, vp, vp, vp, vp, C.c_uint32, vp]
    lib.rtdc_sgd_step_advance_log.restype = i
    lib.rtdc_adam_step.argtypes = [vp, vp, vp, vp, sz, vp, vp]
    lib.rtdc_adam_step_bf16.argtypes = [vp, vp, vp, vp, vp, sz, vp, vp]
    lib.rtdc_adam_step_bf16.restype = i
    lib.rtdc_adam_advance.argtypes = [vp, vp]
    lib.rtdc_counter_advance.argtypes = [vp, vp, vp]
    for name in ("rtdc_gemm", "rtdc_gemm_plan", "rtdc_softmax_ce", "rtdc_colsum", "rtdc_u8_normalize", "rtdc_u8_affine",
                 "rtdc_fold_input_bias",
                 "rtdc_sgd_step", "rtdc_adam_step", "rtdc_adam_advance", "rtdc_counter_advance"):
        getattr(lib, name).restype = i
    # optional groups (declared by their own modules once present)
    return lib


def lib_path() -> Path:
    return _build.LIB_PATH


def load(build_if_missing: bool = True):
    """Return the loaded shared library, building it first when it is absent or stale."""
    global _lib
    if _lib is not None:
        return _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not _build.LIB_PATH.exists() or (build_if_missing and not _build.is_current()):
            if not build_if_missing:
                raise RtdcError(f"{_build.LIB_PATH} is missing; run `python -m ray_torch_distributed_checkpoint_b200.ops._build`")
            try:
                _build.build()
            except Exception as e:  # stale-but-present library is still usable if nvcc is absent
                if not _build.LIB_PATH.exists():
                    raise RtdcError(f"cannot build librtdc_b200.so: {e}") from e
        _lib = _declare(C.CDLL(str(_build.LIB_PATH), mode=C.RTLD_GLOBAL))
        return _lib


def check(rc: int):
    if rc != 0:
        msg = load().rtdc_last_error().decode(errors="replace")
        raise RtdcError(f"librtdc_b200 call failed (code {rc}): {msg}")


def stream_ptr(stream=None) -> int:
    import torch
    s = stream if stream is not None else torch.cuda.current_stream()
    return int(s.cuda_stream)
