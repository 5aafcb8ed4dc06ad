// bf16 LayerNorm (forward; backward with the weight / bias gradients in the same pass) and bf16 column sums (bias gradient
// of a Linear layer) for the GPT-2 / ResNet-50 configurations of BASELINE.json. Not a reference capability (the reference MLP
// has no normalisation layer). Why own kernels: in a GPT-2 small step on a B200 the stock LayerNorm backward spends 86 us per
// call on the weight / bias gradients alone (a [16 384, 768] bf16 input: 50 MB, 0.6 TB/s) plus 26 us on the input gradient,
// 39 us on the forward, and each bias gradient is another strided reduction: together a quarter of the step's non-GEMM time.
// Here a warp owns a row (all of it in registers, fp32 math), so the forward reads x once and writes y once, and the backward
// reads dy and x once, writes dx once and keeps the per-column sums of dy and dy * xhat in registers across the rows a warp
// visits; per-block partial sums are added in block order by a second small kernel (reproducible).
#include <cuda_bf16.h>
#include <stdint.h>

#include "common.h"
#include "../../include/rtdc_b200.h"

namespace rtdc {

constexpr int LN_FWD_THREADS = 256;
constexpr int LN_BWD_THREADS = 256;          // 8 warps, one block per SM (the row in flight + the prefetched one + 48 column sums per lane)
constexpr int LN_MAX_NVL = 4;                // vectors of 8 per lane: C <= 1024

__device__ __forceinline__ void ln_unpack8(const uint4& q, float* x) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&q);
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = __bfloat1622float2(h[i]);
    x[2 * i] = f.x;
    x[2 * i + 1] = f.y;
  }
}
__device__ __forceinline__ uint4 ln_pack8(const float* x) {
  uint4 o;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&o);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(x[2 * i], x[2 * i + 1]);
  return o;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// y = (x - mean) * rstd * gamma + beta; mean / rstd [rows] saved for the backward pass. One warp per row.
template <int NVL>
__global__ void __launch_bounds__(LN_FWD_THREADS)
ln_fwd_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, const __nv_bfloat16* __restrict__ gamma,
              const __nv_bfloat16* __restrict__ beta, float eps, int rows, int C, __nv_bfloat16* __restrict__ y, long long ldy,
              float* __restrict__ mean_out, float* __restrict__ rstd_out) {
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nvec = C >> 3;
  float g[NVL][8], b[NVL][8];
#pragma unroll
  for (int k = 0; k < NVL; ++k) {
    const int v = lane + 32 * k;
    if (v < nvec) {
      ln_unpack8(__ldg(reinterpret_cast<const uint4*>(gamma) + v), g[k]);
      if (beta != nullptr) ln_unpack8(__ldg(reinterpret_cast<const uint4*>(beta) + v), b[k]);
      else {
#pragma unroll
        for (int j = 0; j < 8; ++j) b[k][j] = 0.f;
      }
    }
  }
  const float inv_c = 1.f / static_cast<float>(C);
  const int wpb = LN_FWD_THREADS / 32;
  // software pipeline: the next row's loads are issued before this row's arithmetic, so a warp always has a row in flight
  const int stride = gridDim.x * wpb;
  int row = blockIdx.x * wpb + warp;
  uint4 q[NVL];
  if (row < rows) {
    const uint4* xr = reinterpret_cast<const uint4*>(x + static_cast<long long>(row) * ldx);
#pragma unroll
    for (int k = 0; k < NVL; ++k)
      if (lane + 32 * k < nvec) q[k] = xr[lane + 32 * k];
  }
  for (; row < rows; row += stride) {
    uint4 qn[NVL];
    if (row + stride < rows) {
      const uint4* xn = reinterpret_cast<const uint4*>(x + static_cast<long long>(row + stride) * ldx);
#pragma unroll
      for (int k = 0; k < NVL; ++k)
        if (lane + 32 * k < nvec) qn[k] = xn[lane + 32 * k];
    }
    float xv[NVL][8];
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < NVL; ++k) {
      if (lane + 32 * k < nvec) {
        ln_unpack8(q[k], xv[k]);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += xv[k][j];
      }
    }
    const float mean = warp_sum(s) * inv_c;
    float sq = 0.f;
#pragma unroll
    for (int k = 0; k < NVL; ++k) {
      if (lane + 32 * k < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          const float d = xv[k][j] - mean;
          sq = fmaf(d, d, sq);
        }
      }
    }
    const float rstd = rsqrtf(warp_sum(sq) * inv_c + eps);
    uint4* yr = reinterpret_cast<uint4*>(y + static_cast<long long>(row) * ldy);
#pragma unroll
    for (int k = 0; k < NVL; ++k) {
      const int v = lane + 32 * k;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = fmaf((xv[k][j] - mean) * rstd, g[k][j], b[k][j]);
        yr[v] = ln_pack8(o);
      }
    }
    if (lane == 0) {
      mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
#pragma unroll
    for (int k = 0; k < NVL; ++k) q[k] = qn[k];
  }
}

// dx = rstd * (g - mean_c(g) - xhat * mean_c(g * xhat)), g = dy * gamma, xhat = (x - mean) * rstd;
// partial[block][0][c] = sum over the block's rows of dy * xhat (weight gradient), partial[block][1][c] = sum of dy (bias).
template <int NVL>
__global__ void __launch_bounds__(LN_BWD_THREADS, 1)
ln_bwd_kernel(const __nv_bfloat16* __restrict__ dy, long long lddy, const __nv_bfloat16* __restrict__ x, long long ldx,
              const __nv_bfloat16* __restrict__ gamma, const float* __restrict__ mean, const float* __restrict__ rstd, int rows,
              int C, __nv_bfloat16* __restrict__ dx, long long lddx, float* __restrict__ partial) {
  extern __shared__ __align__(16) float ln_acc[];          // [2][C]
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nvec = C >> 3;
  constexpr int WPB = LN_BWD_THREADS / 32;
  float gm[NVL][8], dg[NVL][8], db[NVL][8];
#pragma unroll
  for (int k = 0; k < NVL; ++k) {
    const int v = lane + 32 * k;
#pragma unroll
    for (int j = 0; j < 8; ++j) { dg[k][j] = 0.f; db[k][j] = 0.f; gm[k][j] = 0.f; }
    if (v < nvec) ln_unpack8(__ldg(reinterpret_cast<const uint4*>(gamma) + v), gm[k]);
  }
  for (int i = threadIdx.x; i < 2 * C; i += LN_BWD_THREADS) ln_acc[i] = 0.f;
  const float inv_c = 1.f / static_cast<float>(C);
  // software pipeline: the next row's loads are issued before this row's arithmetic
  const int stride = gridDim.x * WPB;
  int row = blockIdx.x * WPB + warp;
  uint4 qx[NVL], qd[NVL];
  float mu = 0.f, rs = 0.f;
  if (row < rows) {
    mu = mean[row];
    rs = rstd[row];
    const uint4* xr = reinterpret_cast<const uint4*>(x + static_cast<long long>(row) * ldx);
    const uint4* dyr = reinterpret_cast<const uint4*>(dy + static_cast<long long>(row) * lddy);
#pragma unroll
    for (int k = 0; k < NVL; ++k) {
      const int v = lane + 32 * k;
      if (v < nvec) { qx[k] = xr[v]; qd[k] = dyr[v]; }
    }
  }
  for (; row < rows; row += stride) {
    uint4 nx[NVL], nd[NVL];
    float nmu = 0.f, nrs = 0.f;
    if (row + stride < rows) {
      nmu = mean[row + stride];
      nrs = rstd[row + stride];
      const uint4* xn = reinterpret_cast<const uint4*>(x + static_cast<long long>(row + stride) * ldx);
      const uint4* dyn = reinterpret_cast<const uint4*>(dy + static_cast<long long>(row + stride) * lddy);
#pragma unroll
      for (int k = 0; k < NVL; ++k) {
        const int v = lane + 32 * k;
        if (v < nvec) { nx[k] = xn[v]; nd[k] = dyn[v]; }
      }
    }
    float xh[NVL][8], gv[NVL][8];
    float c1 = 0.f, c2 = 0.f;
#pragma unroll
    for (int k = 0; k < NVL; ++k) {
      if (lane + 32 * k < nvec) {
        float dv[8];
        ln_unpack8(qx[k], xh[k]);
        ln_unpack8(qd[k], dv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          xh[k][j] = (xh[k][j] - mu) * rs;
          gv[k][j] = dv[j] * gm[k][j];
          c1 = fmaf(gv[k][j], xh[k][j], c1);
          c2 += gv[k][j];
          dg[k][j] = fmaf(dv[j], xh[k][j], dg[k][j]);
          db[k][j] += dv[j];
        }
      }
    }
    c1 = warp_sum(c1) * inv_c;
    c2 = warp_sum(c2) * inv_c;
    uint4* dxr = reinterpret_cast<uint4*>(dx + static_cast<long long>(row) * lddx);
#pragma unroll
    for (int k = 0; k < NVL; ++k) {
      const int v = lane + 32 * k;
      if (v < nvec) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rs * (gv[k][j] - c2 - xh[k][j] * c1);
        dxr[v] = ln_pack8(o);
      }
    }
#pragma unroll
    for (int k = 0; k < NVL; ++k) { qx[k] = nx[k]; qd[k] = nd[k]; }
    mu = nmu;
    rs = nrs;
  }
  // the block's warps add their column sums in warp order
  __syncthreads();
  for (int w = 0; w < WPB; ++w) {
    if (warp == w) {
#pragma unroll
      for (int k = 0; k < NVL; ++k) {
        const int v = lane + 32 * k;
        if (v < nvec) {
#pragma unroll
          for (int j = 0; j < 8; ++j) {
            ln_acc[8 * v + j] += dg[k][j];
            ln_acc[C + 8 * v + j] += db[k][j];
          }
        }
      }
    }
    __syncthreads();
  }
  float* dst = partial + static_cast<size_t>(blockIdx.x) * 2 * C;
  for (int i = threadIdx.x; i < 2 * C; i += LN_BWD_THREADS) dst[i] = ln_acc[i];
}

// out[c] = sum over b < nblocks of partial[b * ld + c], c < n; eight warps each add a strided eighth of the partials
// (loads in flight together), then the eight subtotals are added in warp order. out is bf16 or fp32.
__global__ void __launch_bounds__(256)
partial_reduce_kernel(const float* __restrict__ partial, int nblocks, long long ld, int n, void* __restrict__ out, int out_fp32) {
  __shared__ float s_part[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float acc = 0.f;
  if (c < n) {
    for (int b0 = warp; b0 < nblocks; b0 += 64) {
      float v[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int b = b0 + 8 * u;
        v[u] = b < nblocks ? __ldcg(partial + static_cast<long long>(b) * ld + c) : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) acc += v[u];
    }
  }
  s_part[warp][lane] = acc;
  __syncthreads();
  if (warp == 0 && c < n) {
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) tot += s_part[w][lane];
    if (out_fp32) static_cast<float*>(out)[c] = tot;
    else static_cast<__nv_bfloat16*>(out)[c] = __float2bfloat16_rn(tot);
  }
}

// Column sums of a bf16 matrix: block = 32 column vectors (256 columns) x 8 row lanes over one slab of rows.
__global__ void __launch_bounds__(256)
colsum_bf16_kernel(const __nv_bfloat16* __restrict__ x, long long ldx, int rows, int cols, int rows_per_slab,
                   float* __restrict__ partial, long long ldp) {
  __shared__ float red[8][32][9];                           // padded: the final read walks ty for a fixed (tx, j)
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int col0 = (blockIdx.x * 32 + tx) * 8;
  const bool live = col0 + 8 <= cols;
  const int r0 = blockIdx.y * rows_per_slab;
  const int r1 = min(rows, r0 + rows_per_slab);
  float acc[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) acc[j] = 0.f;
  if (live) {
    const __nv_bfloat16* base = x + col0;
    int r = r0 + ty;
    for (; r + 24 < r1; r += 32) {                           // four row loads in flight per thread
      uint4 q[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) q[u] = *reinterpret_cast<const uint4*>(base + static_cast<long long>(r + 8 * u) * ldx);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        float f[8];
        ln_unpack8(q[u], f);
#pragma unroll
        for (int j = 0; j < 8; ++j) acc[j] += f[j];
      }
    }
    for (; r < r1; r += 8) {
      float f[8];
      ln_unpack8(*reinterpret_cast<const uint4*>(base + static_cast<long long>(r) * ldx), f);
#pragma unroll
      for (int j = 0; j < 8; ++j) acc[j] += f[j];
    }
  }
#pragma unroll
  for (int j = 0; j < 8; ++j) red[ty][tx][j] = acc[j];
  __syncthreads();
  // 256 threads finish 256 columns: thread t owns column (t >> 3) * 8 + (t & 7) of the block
  const int cx = threadIdx.x >> 3, cj = threadIdx.x & 7;
  const int col = (blockIdx.x * 32 + cx) * 8 + cj;
  if (col < cols) {
    float tot = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) tot += red[w][cx][cj];
    partial[static_cast<long long>(blockIdx.y) * ldp + col] = tot;
  }
}

static int sm_count_ln() {
  int sms = 0, dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  return sms;
}

static int ln_check(const void* a, long long lda, int rows, int C, const char* what) {
  if (rows <= 0 || C <= 0 || (C & 7) || C > 256 * LN_MAX_NVL) return fail(RTDC_ERR_LIMIT, "%s: the normalised width must be a multiple of 8 and <= %d (got %d)", what, 256 * LN_MAX_NVL, C);
  if ((reinterpret_cast<uintptr_t>(a) & 15u) || (lda & 7) || lda < C) return fail(RTDC_ERR_ARG, "%s: rows must be 16-byte aligned (pitch a multiple of 8 elements, >= C)", what);
  return RTDC_OK;
}

}  // namespace rtdc

using namespace rtdc;

RTDC_API int rtdc_layernorm_fwd_bf16(const void* x, long long ldx, const void* gamma, const void* beta, float eps, int rows, int C,
                                     void* y, long long ldy, float* mean, float* rstd, void* stream) {
  if (!x || !gamma || !y || !mean || !rstd) return fail(RTDC_ERR_ARG, "layernorm_fwd: null pointer");
  int rc = ln_check(x, ldx, rows, C, "layernorm_fwd");
  if (rc == RTDC_OK) rc = ln_check(y, ldy, rows, C, "layernorm_fwd");
  if (rc != RTDC_OK) return rc;
  if ((reinterpret_cast<uintptr_t>(gamma) & 15u) || (beta && (reinterpret_cast<uintptr_t>(beta) & 15u)))
    return fail(RTDC_ERR_ARG, "layernorm_fwd: weight / bias must be 16-byte aligned");
  const int wpb = LN_FWD_THREADS / 32;
  const int want = ceil_div(rows, wpb), cap = sm_count_ln() * 8;
  const int grid = want < cap ? want : cap;
  const int nvl = ceil_div(C >> 3, 32);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* gp = static_cast<const __nv_bfloat16*>(gamma);
  const __nv_bfloat16* bp = static_cast<const __nv_bfloat16*>(beta);
  __nv_bfloat16* yp = static_cast<__nv_bfloat16*>(y);
  switch (nvl) {
    case 1: ln_fwd_kernel<1><<<grid, LN_FWD_THREADS, 0, st>>>(xp, ldx, gp, bp, eps, rows, C, yp, ldy, mean, rstd); break;
    case 2: ln_fwd_kernel<2><<<grid, LN_FWD_THREADS, 0, st>>>(xp, ldx, gp, bp, eps, rows, C, yp, ldy, mean, rstd); break;
    case 3: ln_fwd_kernel<3><<<grid, LN_FWD_THREADS, 0, st>>>(xp, ldx, gp, bp, eps, rows, C, yp, ldy, mean, rstd); break;
    default: ln_fwd_kernel<4><<<grid, LN_FWD_THREADS, 0, st>>>(xp, ldx, gp, bp, eps, rows, C, yp, ldy, mean, rstd); break;
  }
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API size_t rtdc_layernorm_bwd_scratch_bytes(int rows, int C) {
  (void)rows;
  return static_cast<size_t>(sm_count_ln()) * 2 * static_cast<size_t>(C) * sizeof(float);
}

RTDC_API int rtdc_layernorm_bwd_bf16(const void* dy, long long lddy, const void* x, long long ldx, const void* gamma, const float* mean,
                                     const float* rstd, int rows, int C, void* dx, long long lddx, void* dgamma, void* dbeta,
                                     void* scratch, size_t scratch_bytes, void* stream) {
  if (!dy || !x || !gamma || !mean || !rstd || !dx || !dgamma || !scratch) return fail(RTDC_ERR_ARG, "layernorm_bwd: null pointer");
  int rc = ln_check(dy, lddy, rows, C, "layernorm_bwd");
  if (rc == RTDC_OK) rc = ln_check(x, ldx, rows, C, "layernorm_bwd");
  if (rc == RTDC_OK) rc = ln_check(dx, lddx, rows, C, "layernorm_bwd");
  if (rc != RTDC_OK) return rc;
  if (reinterpret_cast<uintptr_t>(gamma) & 15u) return fail(RTDC_ERR_ARG, "layernorm_bwd: weight must be 16-byte aligned");
  constexpr int WPB = LN_BWD_THREADS / 32;
  const int want = ceil_div(rows, WPB), cap = sm_count_ln();
  const int grid = want < cap ? want : cap;
  if (scratch_bytes < static_cast<size_t>(grid) * 2 * C * sizeof(float))
    return fail(RTDC_ERR_ARG, "layernorm_bwd needs %zu bytes of scratch (got %zu)", rtdc_layernorm_bwd_scratch_bytes(rows, C), scratch_bytes);
  float* partial = static_cast<float*>(scratch);
  const size_t smem = 2 * static_cast<size_t>(C) * sizeof(float);
  const int nvl = ceil_div(C >> 3, 32);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const __nv_bfloat16* dyp = static_cast<const __nv_bfloat16*>(dy);
  const __nv_bfloat16* xp = static_cast<const __nv_bfloat16*>(x);
  const __nv_bfloat16* gp = static_cast<const __nv_bfloat16*>(gamma);
  __nv_bfloat16* dxp = static_cast<__nv_bfloat16*>(dx);
  switch (nvl) {
    case 1: ln_bwd_kernel<1><<<grid, LN_BWD_THREADS, smem, st>>>(dyp, lddy, xp, ldx, gp, mean, rstd, rows, C, dxp, lddx, partial); break;
    case 2: ln_bwd_kernel<2><<<grid, LN_BWD_THREADS, smem, st>>>(dyp, lddy, xp, ldx, gp, mean, rstd, rows, C, dxp, lddx, partial); break;
    case 3: ln_bwd_kernel<3><<<grid, LN_BWD_THREADS, smem, st>>>(dyp, lddy, xp, ldx, gp, mean, rstd, rows, C, dxp, lddx, partial); break;
    default: ln_bwd_kernel<4><<<grid, LN_BWD_THREADS, smem, st>>>(dyp, lddy, xp, ldx, gp, mean, rstd, rows, C, dxp, lddx, partial); break;
  }
  RTDC_CUDA_CHECK(cudaGetLastError());
  // weight gradient = columns [0, C) of the partials, bias gradient = columns [C, 2C)
  partial_reduce_kernel<<<ceil_div(C, 32), 256, 0, st>>>(partial, grid, 2LL * C, C, dgamma, 0);
  if (dbeta) partial_reduce_kernel<<<ceil_div(C, 32), 256, 0, st>>>(partial + C, grid, 2LL * C, C, dbeta, 0);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

static void colsum_bf16_plan(int rows, int cols, int* gx, int* slabs, int* rows_per_slab) {
  *gx = ceil_div(cols, 256);
  int s = ceil_div(2 * sm_count_ln(), *gx);
  const int max_s = ceil_div(rows, 64);                     // at least 64 rows per slab
  if (s > max_s) s = max_s;
  if (s < 1) s = 1;
  *rows_per_slab = round_up(ceil_div(rows, s), 8);
  *slabs = ceil_div(rows, *rows_per_slab);
}

RTDC_API size_t rtdc_colsum_bf16_scratch_bytes(int rows, int cols) {
  int gx, slabs, rps;
  colsum_bf16_plan(rows, cols, &gx, &slabs, &rps);
  return static_cast<size_t>(slabs) * static_cast<size_t>(cols) * sizeof(float);
}

RTDC_API int rtdc_colsum_bf16(const void* x, long long ldx, int rows, int cols, void* out, int out_fp32, void* scratch,
                              size_t scratch_bytes, void* stream) {
  if (!x || !out || !scratch) return fail(RTDC_ERR_ARG, "colsum_bf16: null pointer");
  if (rows <= 0 || cols <= 0 || (cols & 7) || (ldx & 7) || ldx < cols || (reinterpret_cast<uintptr_t>(x) & 15u))
    return fail(RTDC_ERR_ARG, "colsum_bf16: cols and the row pitch must be multiples of 8, rows 16-byte aligned");
  int gx, slabs, rps;
  colsum_bf16_plan(rows, cols, &gx, &slabs, &rps);
  if (scratch_bytes < static_cast<size_t>(slabs) * cols * sizeof(float))
    return fail(RTDC_ERR_ARG, "colsum_bf16 needs %zu bytes of scratch (got %zu)", rtdc_colsum_bf16_scratch_bytes(rows, cols), scratch_bytes);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  float* partial = static_cast<float*>(scratch);
  colsum_bf16_kernel<<<dim3(gx, slabs), 256, 0, st>>>(static_cast<const __nv_bfloat16*>(x), ldx, rows, cols, rps, partial, cols);
  RTDC_CUDA_CHECK(cudaGetLastError());
  partial_reduce_kernel<<<ceil_div(cols, 32), 256, 0, st>>>(partial, slabs, cols, cols, out, out_fp32);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}
