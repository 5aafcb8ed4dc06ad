// Softmax cross-entropy over a wide vocabulary fused with its own backward, for the tied LM head of the GPT-2
// configuration (BASELINE.json config #5; not a reference capability: the reference's only loss is the 10-class
// nn.CrossEntropyLoss of my_ray_module.py:141, which csrc/mlp_head.cu and softmax_ce_kernel serve).
//
// Stock PyTorch runs `F.cross_entropy(logits.float(), targets)` on the bf16 logits of the head GEMM as a chain of
// full-size fp32 tensors: the upcast, log_softmax and its saved output, the gradient of each on the way back —
// about 30 GB of HBM traffic for 16 x 1024 tokens x 50 257 classes, more than the rest of the step's elementwise work
// together. Here ONE pass over the bf16 logits does all of it: a block owns a row, reads it once (16-byte vectors, kept
// in shared memory), reduces max / sum-of-exponentials online in fp32, writes the row's loss and overwrites the logits
// IN PLACE with d loss / d logits = (softmax - onehot) * grad_scale in bf16 — the operand the head's dgrad / wgrad GEMMs
// (csrc/gemm_bf16.cu) then read as stored. HBM traffic: the logits once in, once out.
#include <cuda_bf16.h>
#include <math.h>
#include <stdint.h>

#include "common.h"
#include "../../include/rtdc_b200.h"

namespace rtdc {

constexpr int LM_THREADS = 512;
constexpr int LM_WARPS = LM_THREADS / 32;
constexpr int LM_BLOCKS_PER_SM = 2;
constexpr int LM_STAGE_MAX_BYTES = 110 * 1024;      // per block; two blocks per SM
constexpr float LM_LOG2E = 1.4426950408889634f;
constexpr float LM_LN2 = 0.6931471805599453f;

struct LmCeParams {
  __nv_bfloat16* z;            // [rows, ld] in: logits, out: dlogits
  long long ld;
  const long long* targets;    // [rows]
  int rows, V;
  long long ignore_index;
  const float* grad_scale;     // device scalar (1 / number of counted rows for a mean reduction)
  float* row_loss;             // [rows], natural-log loss of each row (0 for ignored rows)
};

__device__ __forceinline__ void unpack8(const uint4& q, float* x) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(h[i]);
    x[2 * i] = f.x;
    x[2 * i + 1] = f.y;
  }
}

// (m, s) represent s * 2^m; merge two such partial sums of exponentials
__device__ __forceinline__ void lse_merge(float& m, float& s, float m2, float s2) {
  const float mn = fmaxf(m, m2);
  if (mn == -INFINITY) return;                       // both empty
  s = s * exp2f(m - mn) + s2 * exp2f(m2 - mn);
  m = mn;
}

__device__ __forceinline__ void lse_push8(float& m, float& s, const float* x) {
  float xm = x[0];
#pragma unroll
  for (int j = 1; j < 8; ++j) xm = fmaxf(xm, x[j]);
  const float ym = xm * LM_LOG2E;
  if (ym > m) {
    s *= exp2f(m - ym);                              // m == -inf: s is 0 and stays 0
    m = ym;
  }
  float acc = 0.f;
#pragma unroll
  for (int j = 0; j < 8; ++j) acc += exp2f(fmaf(x[j], LM_LOG2E, -m));
  s += acc;
}

template <bool STAGE>
__global__ void __launch_bounds__(LM_THREADS, LM_BLOCKS_PER_SM)
lm_head_ce_kernel(const LmCeParams p) {
  extern __shared__ __align__(16) unsigned char lm_smem[];
  uint4* srow = reinterpret_cast<uint4*>(lm_smem);
  __shared__ float sm_m[LM_WARPS], sm_s[LM_WARPS];
  __shared__ float s_xt;

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int nv = p.V >> 3, tail = p.V & 7;
  const float gscale = __ldg(p.grad_scale);

  for (int row = blockIdx.x; row < p.rows; row += gridDim.x) {
    __nv_bfloat16* zr = p.z + static_cast<long long>(row) * p.ld;
    uint4* zv = reinterpret_cast<uint4*>(zr);
    const long long tgt = p.targets[row];
    const bool counted = tgt != p.ignore_index && tgt >= 0 && tgt < p.V;      // block-uniform
    if (!counted) {
      const uint4 zero = make_uint4(0u, 0u, 0u, 0u);
      for (int i = t; i < nv; i += LM_THREADS) zv[i] = zero;
      if (t < tail) zr[8 * nv + t] = __float2bfloat16_rn(0.f);
      if (t == 0) p.row_loss[row] = 0.f;
      continue;
    }

    // ---- pass 1: the row once from global memory; online max / sum of 2^(x log2e - m)
    float m = -INFINITY, s = 0.f;
    {
      int i = t;
      for (; i + 3 * LM_THREADS < nv; i += 4 * LM_THREADS) {          // four 16-byte loads in flight per thread
        uint4 q[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) q[u] = zv[i + u * LM_THREADS];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (STAGE) srow[i + u * LM_THREADS] = q[u];
          float x[8];
          unpack8(q[u], x);
          lse_push8(m, s, x);
        }
      }
      for (; i < nv; i += LM_THREADS) {
        const uint4 q = zv[i];
        if (STAGE) srow[i] = q;
        float x[8];
        unpack8(q, x);
        lse_push8(m, s, x);
      }
    }
    float x_tail = 0.f;
    if (t < tail) {
      x_tail = __bfloat162float(zr[8 * nv + t]);
      lse_merge(m, s, x_tail * LM_LOG2E, 1.f);
    }
    if (t == 0) s_xt = __bfloat162float(zr[tgt]);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
      const float m2 = __shfl_xor_sync(0xffffffffu, m, o);
      const float s2 = __shfl_xor_sync(0xffffffffu, s, o);
      lse_merge(m, s, m2, s2);
    }
    if (lane == 0) { sm_m[warp] = m; sm_s[warp] = s; }
    __syncthreads();
    float M = sm_m[0], S = sm_s[0];
#pragma unroll
    for (int w = 1; w < LM_WARPS; ++w) lse_merge(M, S, sm_m[w], sm_s[w]);      // same order in every thread: identical M, S
    const float xt = s_xt;
    if (t == 0) p.row_loss[row] = LM_LN2 * (M + log2f(S)) - xt;

    // ---- pass 2: d logits = (softmax - onehot) * grad_scale, in place
    const float inv = gscale / S;
    const int tv = static_cast<int>(tgt >> 3), tj = static_cast<int>(tgt & 7);
    for (int i = t; i < nv; i += LM_THREADS) {
      const uint4 q = STAGE ? srow[i] : zv[i];
      float x[8];
      unpack8(q, x);
#pragma unroll
      for (int j = 0; j < 8; ++j) x[j] = exp2f(fmaf(x[j], LM_LOG2E, -M)) * inv;
      if (i == tv) {
#pragma unroll
        for (int j = 0; j < 8; ++j)
          if (j == tj) x[j] -= gscale;
      }
      uint4 o;
      __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
      for (int j = 0; j < 4; ++j) h[j] = __floats2bfloat162_rn(x[2 * j], x[2 * j + 1]);
      zv[i] = o;
    }
    if (t < tail) {
      float d = exp2f(fmaf(x_tail, LM_LOG2E, -M)) * inv;
      if (8 * nv + t == tgt) d -= gscale;
      zr[8 * nv + t] = __float2bfloat16_rn(d);
    }
    __syncthreads();                                 // sm_m / sm_s / s_xt / srow are reused by the next row
  }
}

}  // namespace rtdc

using namespace rtdc;

RTDC_API int rtdc_lm_head_ce_bf16(void* logits, long long ld, const int64_t* targets, int rows, int vocab,
                                  long long ignore_index, const float* grad_scale, float* row_loss, void* stream) {
  if (!logits || !targets || !grad_scale || !row_loss) return fail(RTDC_ERR_ARG, "lm_head_ce: null pointer");
  if (rows <= 0 || vocab <= 0 || ld < vocab) return fail(RTDC_ERR_ARG, "lm_head_ce: bad shape rows=%d vocab=%d ld=%lld", rows, vocab, ld);
  if ((reinterpret_cast<uintptr_t>(logits) & 15u) || (ld & 7))
    return fail(RTDC_ERR_ARG, "lm_head_ce: logits must be 16-byte aligned with a row pitch that is a multiple of 8 elements");
  LmCeParams p;
  p.z = static_cast<__nv_bfloat16*>(logits); p.ld = ld; p.targets = reinterpret_cast<const long long*>(targets);
  p.rows = rows; p.V = vocab; p.ignore_index = ignore_index; p.grad_scale = grad_scale; p.row_loss = row_loss;
  int sms = 0, dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  const int cap = sms * LM_BLOCKS_PER_SM;
  const int grid = rows < cap ? rows : cap;
  const size_t stage_bytes = static_cast<size_t>(vocab >> 3) * 16;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (stage_bytes <= static_cast<size_t>(LM_STAGE_MAX_BYTES)) {
    RTDC_CUDA_CHECK(cudaFuncSetAttribute(lm_head_ce_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(stage_bytes)));
    lm_head_ce_kernel<true><<<grid, LM_THREADS, stage_bytes, st>>>(p);
  } else {
    lm_head_ce_kernel<false><<<grid, LM_THREADS, 0, st>>>(p);       // wider rows: the second pass re-reads global memory (L2)
  }
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}
