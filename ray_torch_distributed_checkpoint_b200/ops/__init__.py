"""Hand-written sm_100a operators (tcgen05 GEMMs, fused loss/optimizer/normalisation kernels).

Importing this package does not load the shared library; the first op call does (and fails loudly
if ``librtdc_b200.so`` cannot be found or built).
"""
from .gemm import (PRECISION_FP32, PRECISION_TF32, gemm, gemm_plan, linear_dgrad, linear_forward, linear_wgrad,
                   scratch_for)
from .elementwise import (AdamState, SgdState, act_backward, adam_step, colsum, counter_advance, fold_input_bias, mlp_head_train,
                          h2d_pieces, sgd_step, softmax_cross_entropy, u8_affine, u8_gather_rows, u8_normalize)
from .gemm_bf16 import LinearBf16, gemm_bf16, use_bf16_linears
from .lm_head import lm_head_cross_entropy, softmax_ce_bf16_
from .layernorm import LayerNormBf16, colsum_bf16, layer_norm_bf16, use_fused_layernorm
from ._lib import RtdcError, lib_path, load as load_library

__all__ = [
    "PRECISION_FP32", "PRECISION_TF32", "gemm", "gemm_plan", "linear_forward", "linear_dgrad", "linear_wgrad",
    "scratch_for", "act_backward", "softmax_cross_entropy", "mlp_head_train", "colsum", "u8_gather_rows", "h2d_pieces", "u8_normalize", "u8_affine", "fold_input_bias", "sgd_step", "adam_step", "counter_advance",
    "gemm_bf16", "LinearBf16", "use_bf16_linears", "lm_head_cross_entropy", "softmax_ce_bf16_", "LayerNormBf16", "layer_norm_bf16", "use_fused_layernorm", "colsum_bf16", "SgdState", "AdamState", "RtdcError", "lib_path", "load_library",
]
