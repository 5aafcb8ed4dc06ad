// Memory-bound sm_100a kernels around the GEMMs: fused softmax-cross-entropy forward+backward
// (K4/K5), deterministic column sums for bias gradients (K6), uint8->fp32 input normalisation
// (K9 fused with the reference's ToTensor+Normalize), fused multi-tensor SGD / Adam over a flat
// arena (K7). All are 128-bit vectorised grid-stride kernels sized to 148 SMs x resident CTAs.
#include <cuda_bf16.h>

#include "common.h"
#include "ptx.cuh"
#include "../../include/rtdc_b200.h"

namespace rtdc {

constexpr int CE_THREADS = 256;
constexpr int CE_MAX_BLOCKS = 1184;  // 148 SMs x 8 resident CTAs

// ---------------------------------------------------------------------------------------------
// softmax cross-entropy. One thread per row (classes <= 32 keeps a row in registers); block
// partial sums go to scratch and the last block to finish adds them in block order, so the loss
// is bit-reproducible. scratch layout: float partial[CE_MAX_BLOCKS]; u64 correct[CE_MAX_BLOCKS];
// u32 counter.
struct CeScratch {
  unsigned int counter;
  unsigned int pad[63];
  float partial[CE_MAX_BLOCKS];
  unsigned long long correct[CE_MAX_BLOCKS];
  float dbias[CE_MAX_BLOCKS][32];   // per-block column sums of dlogits (bias gradient of the last layer)
};

template <int MAXC>
__global__ void __launch_bounds__(CE_THREADS)
softmax_ce_kernel(const float* __restrict__ logits, int ldz, const long long* __restrict__ labels, int rows,
                  int classes, float* __restrict__ loss_out, float* __restrict__ dlogits, float grad_scale,
                  int relu_mask, unsigned long long* correct_accum, float* loss_sum_accum, float* __restrict__ dbias_out,
                  CeScratch* scratch) {
  __shared__ float s_loss[CE_THREADS / 32];
  __shared__ unsigned int s_correct[CE_THREADS / 32];
  __shared__ float s_db[CE_THREADS / 32][32];
  __shared__ unsigned int s_last;
  float my_loss = 0.f;
  unsigned int my_correct = 0;
  const bool want_db = dbias_out != nullptr && dlogits != nullptr;
  float gsum[MAXC];
#pragma unroll
  for (int j = 0; j < MAXC; ++j) gsum[j] = 0.f;
  for (int row = blockIdx.x * blockDim.x + threadIdx.x; row < rows; row += gridDim.x * blockDim.x) {
    const float* z = logits + static_cast<size_t>(row) * ldz;
    float v[MAXC];
    float mx = -INFINITY;
    int arg = 0;
    if ((ldz & 3) == 0) {
#pragma unroll
      for (int q = 0; q < MAXC / 4; ++q) {
        if (q * 4 < ldz) {
          const float4 t = __ldg(reinterpret_cast<const float4*>(z) + q);
          v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
        }
      }
    } else {
#pragma unroll
      for (int j = 0; j < MAXC; ++j)
        if (j < classes) v[j] = __ldg(z + j);
    }
#pragma unroll
    for (int j = 0; j < MAXC; ++j) {
      if (j < classes && v[j] > mx) { mx = v[j]; arg = j; }  // first maximum, like torch.argmax
    }
    float sum = 0.f;
    float e[MAXC];
#pragma unroll
    for (int j = 0; j < MAXC; ++j) {
      e[j] = (j < classes) ? __expf(v[j] - mx) : 0.f;
      sum += e[j];
    }
    const long long y = labels[row];
    float zy = 0.f;
#pragma unroll
    for (int j = 0; j < MAXC; ++j)
      if (j == y) zy = v[j];
    my_loss += (mx + __logf(sum)) - zy;
    my_correct += (arg == y) ? 1u : 0u;
    if (dlogits != nullptr) {
      const float inv = grad_scale / sum;
      float* dz = dlogits + static_cast<size_t>(row) * ldz;
      float g[MAXC];
#pragma unroll
      for (int j = 0; j < MAXC; ++j) {
        float t = 0.f;
        if (j < classes) {
          t = e[j] * inv - ((j == y) ? grad_scale : 0.f);
          if (relu_mask && !(v[j] > 0.f)) t = 0.f;
        }
        g[j] = t;
        gsum[j] += t;
      }
      if ((ldz & 3) == 0) {
#pragma unroll
        for (int q = 0; q < MAXC / 4; ++q)
          if (q * 4 < ldz) reinterpret_cast<float4*>(dz)[q] = make_float4(g[4 * q], g[4 * q + 1], g[4 * q + 2], g[4 * q + 3]);
      } else {
#pragma unroll
        for (int j = 0; j < MAXC; ++j)
          if (j < ldz) dz[j] = g[j];
      }
    }
  }
  // block reduction
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    my_loss += __shfl_xor_sync(0xffffffffu, my_loss, o);
    my_correct += __shfl_xor_sync(0xffffffffu, my_correct, o);
  }
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (want_db) {
#pragma unroll
    for (int j = 0; j < MAXC; ++j) {
      float t = gsum[j];
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
      if (lane == 0) s_db[warp][j] = t;
    }
  }
  if (lane == 0) { s_loss[warp] = my_loss; s_correct[warp] = my_correct; }
  __syncthreads();
  if (want_db && threadIdx.x < 32) {
    float t = 0.f;
    if (threadIdx.x < MAXC)
      for (int w = 0; w < CE_THREADS / 32; ++w) t += s_db[w][threadIdx.x];
    scratch->dbias[blockIdx.x][threadIdx.x] = t;
  }
  if (threadIdx.x == 0) {
    float bl = 0.f;
    unsigned int bc = 0;
    for (int w = 0; w < CE_THREADS / 32; ++w) { bl += s_loss[w]; bc += s_correct[w]; }
    scratch->partial[blockIdx.x] = bl;
    scratch->correct[blockIdx.x] = bc;
  }
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) {
    const unsigned int old = atomicAdd(&scratch->counter, 1u);
    s_last = (old == gridDim.x - 1) ? 1u : 0u;
  }
  __syncthreads();
  if (s_last) {
    // last block to finish: sum the per-block partials with all 256 threads (independent loads in
    // flight), in an order that depends only on the grid size, so the result is reproducible
    __threadfence();
    __shared__ float f_loss[CE_THREADS];
    __shared__ unsigned long long f_corr[CE_THREADS];
    float tl = 0.f;
    unsigned long long tc = 0;
    for (unsigned int b = threadIdx.x; b < gridDim.x; b += CE_THREADS) {
      tl += __ldcg(&scratch->partial[b]);
      tc += __ldcg(&scratch->correct[b]);
    }
    f_loss[threadIdx.x] = tl;
    f_corr[threadIdx.x] = tc;
    if (want_db) {
      const int c = threadIdx.x & 31, part = threadIdx.x >> 5;
      float t = 0.f;
      for (unsigned int b = part; b < gridDim.x; b += CE_THREADS / 32) t += __ldcg(&scratch->dbias[b][c]);
      s_db[part][c] = t;
    }
    __syncthreads();
    for (int o = CE_THREADS / 2; o > 0; o >>= 1) {
      if (threadIdx.x < o) {
        f_loss[threadIdx.x] += f_loss[threadIdx.x + o];
        f_corr[threadIdx.x] += f_corr[threadIdx.x + o];
      }
      __syncthreads();
    }
    if (want_db && threadIdx.x < classes && threadIdx.x < 32) {
      float t = 0.f;
      for (int w = 0; w < CE_THREADS / 32; ++w) t += s_db[w][threadIdx.x];
      dbias_out[threadIdx.x] = t;
    }
    if (threadIdx.x == 0) {
      const float total = f_loss[0];
      if (loss_out) *loss_out = total / static_cast<float>(rows);
      if (loss_sum_accum) *loss_sum_accum += total / static_cast<float>(rows);  // per-batch mean, summed (ref :168)
      if (correct_accum) *correct_accum += f_corr[0];
      scratch->counter = 0u;  // ready for the next launch
    }
  }
}

// ---------------------------------------------------------------------------------------------
// column sums: stage 1 = each block sums a row slab for 128 columns into partial[slab][col];
// last block per column group sums the slabs in order.
constexpr int CS_THREADS = 256;
constexpr int CS_COLS = 128;     // columns per block: 32 lanes x float4
constexpr int CS_MAX_SLABS = 64;   // (the built-in scratch of rtdc_colsum_noalloc is sized for this)

__global__ void __launch_bounds__(CS_THREADS)
colsum_kernel(const float* __restrict__ x, int ld, int rows, int cols, float* __restrict__ out, float* partial,
              unsigned int* counters, int rows_per_slab) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float4 s_acc[CS_THREADS / 32][32];
  __shared__ unsigned int s_last;
  const int cg = blockIdx.x, slab = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c0 = cg * CS_COLS + lane * 4;
  const int r_begin = slab * rows_per_slab, r_end = min(rows, r_begin + rows_per_slab);
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const bool vec = ((ld & 3) == 0) && (c0 + 3 < cols) && ((reinterpret_cast<uintptr_t>(x) & 15u) == 0);
  constexpr int ROWS_ILP = 8;   // independent row loads in flight per thread (the loop is latency bound otherwise)
  constexpr int WARPS = CS_THREADS / 32;
  int r = r_begin + warp;
  if (vec) {
    // every batch (the last, partial one included) keeps ROWS_ILP independent loads in flight
    for (; r < r_end; r += ROWS_ILP * WARPS) {
      float4 t[ROWS_ILP];
#pragma unroll
      for (int u = 0; u < ROWS_ILP; ++u) {
        const int rr = r + u * WARPS;
        t[u] = rr < r_end ? __ldg(reinterpret_cast<const float4*>(x + static_cast<size_t>(rr) * ld + c0)) : make_float4(0.f, 0.f, 0.f, 0.f);
      }
#pragma unroll
      for (int u = 0; u < ROWS_ILP; ++u) { acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w; }
    }
  }
  for (; r < r_end; r += WARPS) {
    const float* src = x + static_cast<size_t>(r) * ld + c0;
    if (vec) {
      const float4 t = __ldg(reinterpret_cast<const float4*>(src));
      acc.x += t.x; acc.y += t.y; acc.z += t.z; acc.w += t.w;
    } else {
      if (c0 < cols) acc.x += __ldg(src);
      if (c0 + 1 < cols) acc.y += __ldg(src + 1);
      if (c0 + 2 < cols) acc.z += __ldg(src + 2);
      if (c0 + 3 < cols) acc.w += __ldg(src + 3);
    }
  }
  s_acc[warp][lane] = acc;
  __syncthreads();
  if (warp == 0) {
    float4 t = s_acc[0][lane];
    for (int w = 1; w < WARPS; ++w) {
      const float4 u = s_acc[w][lane];
      t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
    }
    float* dst = partial + (static_cast<size_t>(slab) * gridDim.x + cg) * CS_COLS + lane * 4;
    *reinterpret_cast<float4*>(dst) = t;
    __threadfence();
    __syncwarp();
    if (lane == 0) {
      const unsigned int old = atomicAdd(&counters[cg], 1u);
      s_last = (old == gridDim.y - 1) ? 1u : 0u;
      if (s_last) counters[cg] = 0u;
    }
  }
  __syncthreads();
  if (s_last) {
    // last block of this column group: every warp sums a strided subset of the slabs (loads independent),
    // then warp 0 adds the eight subtotals in warp order -> same result for the same grid, every time
    __threadfence();
    float4 tot = make_float4(0.f, 0.f, 0.f, 0.f);
    for (unsigned int sl = warp; sl < gridDim.y; sl += WARPS) {
      const float4 u = __ldcg(reinterpret_cast<const float4*>(partial + (static_cast<size_t>(sl) * gridDim.x + cg) * CS_COLS + lane * 4));
      tot.x += u.x; tot.y += u.y; tot.z += u.z; tot.w += u.w;
    }
    __syncthreads();
    s_acc[warp][lane] = tot;
    __syncthreads();
    if (warp == 0) {
      float4 t = s_acc[0][lane];
      for (int w = 1; w < WARPS; ++w) {
        const float4 u = s_acc[w][lane];
        t.x += u.x; t.y += u.y; t.z += u.z; t.w += u.w;
      }
      if (c0 < cols) out[c0] = t.x;
      if (c0 + 1 < cols) out[c0 + 1] = t.y;
      if (c0 + 2 < cols) out[c0 + 2] = t.z;
      if (c0 + 3 < cols) out[c0 + 3] = t.w;
    }
  }
}

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
u8_normalize_kernel(const uint8_t* __restrict__ src, float* __restrict__ dst, size_t n, float scale, float shift) {
  // 16 pixels per thread per iteration: one 128-bit load, four 128-bit stores
  const size_t nvec = n / 16;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i0 = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i0 < nvec; i0 += 2 * stride) {
    const size_t idx[2] = {i0, i0 + stride};
    uint4 w[2];
#pragma unroll
    for (int u = 0; u < 2; ++u)
      if (idx[u] < nvec) w[u] = __ldg(reinterpret_cast<const uint4*>(src) + idx[u]);
#pragma unroll
    for (int u = 0; u < 2; ++u) {
      if (idx[u] >= nvec) continue;
      const uint32_t ws[4] = {w[u].x, w[u].y, w[u].z, w[u].w};
      float4* o = reinterpret_cast<float4*>(dst) + idx[u] * 4;
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        float4 f;
        f.x = static_cast<float>(ws[q] & 0xFFu) * scale + shift;
        f.y = static_cast<float>((ws[q] >> 8) & 0xFFu) * scale + shift;
        f.z = static_cast<float>((ws[q] >> 16) & 0xFFu) * scale + shift;
        f.w = static_cast<float>(ws[q] >> 24) * scale + shift;
        o[q] = f;
      }
    }
  }
  if (blockIdx.x == 0) {
    for (size_t i = nvec * 16 + threadIdx.x; i < n; i += blockDim.x) dst[i] = static_cast<float>(src[i]) * scale + shift;
  }
}

// ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
sgd_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, size_t n,
           rtdc_sgd_hyper* hp, uint32_t* step_counter, int advance, const float* loss, float* loss_log, uint32_t log_mask) {
  pdl_launch_dependents();
  pdl_wait();
  const rtdc_sgd_hyper h = *hp;
  const size_t nvec = n / 4;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  const bool use_m = h.momentum != 0.f;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + i);
    float4 mv = use_m ? reinterpret_cast<float4*>(m)[i] : make_float4(0.f, 0.f, 0.f, 0.f);
    float* pp = &pv.x; const float* gg = &gv.x; float* mm = &mv.x;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      float gr = gg[j] * h.grad_scale + h.weight_decay * pp[j];
      if (use_m) {
        mm[j] = h.first_step ? gr : h.momentum * mm[j] + (1.f - h.dampening) * gr;
        gr = h.nesterov ? gr + h.momentum * mm[j] : mm[j];
      }
      pp[j] -= h.lr * gr;
    }
    reinterpret_cast<float4*>(p)[i] = pv;
    if (use_m) reinterpret_cast<float4*>(m)[i] = mv;
  }
  if (blockIdx.x == 0) {
    for (size_t i = nvec * 4 + threadIdx.x; i < n; i += blockDim.x) {
      float gr = g[i] * h.grad_scale + h.weight_decay * p[i];
      if (use_m) {
        const float mi = h.first_step ? gr : h.momentum * m[i] + (1.f - h.dampening) * gr;
        m[i] = mi;
        gr = h.nesterov ? gr + h.momentum * mi : mi;
      }
      p[i] -= h.lr * gr;
    }
  }
  if (advance) {
    // the block that finishes last clears first_step and bumps the dropout step counter (what a separate one-thread kernel
    // used to do): every block has read its copy of the hyper-parameters by then
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      const int ticket = atomicAdd(&hp->pad, 1);
      if (ticket == static_cast<int>(gridDim.x) - 1) {
        hp->first_step = 0;
        hp->pad = 0;
        if (step_counter != nullptr) {
          const uint32_t c = step_counter[0];
          if (loss_log != nullptr && loss != nullptr) loss_log[c & log_mask] = *loss;   // pinned host ring: this step's loss
          step_counter[0] = c + 1u;
        }
        __threadfence();
      }
    }
  }
}

__device__ __forceinline__ void adam_update(float& p, float g, float& m, float& v, const rtdc_adam_hyper& h, float bc1,
                                            float bc2_sqrt) {
  g *= h.grad_scale;
  if (h.decoupled) p *= (1.f - h.lr * h.weight_decay);
  else g += h.weight_decay * p;
  m = h.beta1 * m + (1.f - h.beta1) * g;
  v = h.beta2 * v + (1.f - h.beta2) * g * g;
  const float denom = sqrtf(v) / bc2_sqrt + h.eps;
  p -= (h.lr / bc1) * (m / denom);
}

__global__ void __launch_bounds__(256)
adam_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m, float* __restrict__ v, size_t n,
            const rtdc_adam_hyper* __restrict__ hp) {
  const rtdc_adam_hyper h = *hp;
  const float t = static_cast<float>(h.step);
  const float bc1 = 1.f - powf(h.beta1, t);
  const float bc2_sqrt = sqrtf(1.f - powf(h.beta2, t));
  const size_t nvec = n / 4;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = __ldg(reinterpret_cast<const float4*>(g) + i);
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    adam_update(pv.x, gv.x, mv.x, vv.x, h, bc1, bc2_sqrt);
    adam_update(pv.y, gv.y, mv.y, vv.y, h, bc1, bc2_sqrt);
    adam_update(pv.z, gv.z, mv.z, vv.z, h, bc1, bc2_sqrt);
    adam_update(pv.w, gv.w, mv.w, vv.w, h, bc1, bc2_sqrt);
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
  }
  if (blockIdx.x == 0) {
    for (size_t i = nvec * 4 + threadIdx.x; i < n; i += blockDim.x) adam_update(p[i], g[i], m[i], v[i], h, bc1, bc2_sqrt);
  }
}

// dz[r, c] = dy[r, c] * (y[r, c] > 0 ? scale : 0) for c < cols, 0 for cols <= c < cols_out: the ReLU(+dropout) backward of a
// layer whose gradient does not come out of a fused dgrad epilogue (the last layer of a chain), written with the padded row
// pitch the TMA-fed wgrad/dgrad GEMMs need.
__global__ void __launch_bounds__(256)
act_backward_kernel(const float* __restrict__ dy, int lddy, const float* __restrict__ y, int ldy, int rows, int cols, int cols_out,
                    float scale, float* __restrict__ dz, int lddz) {
  const size_t total = static_cast<size_t>(rows) * cols_out;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / cols_out), c = static_cast<int>(i % cols_out);
    float v = 0.f;
    if (c < cols) {
      v = dy[static_cast<size_t>(r) * lddy + c];
      if (y != nullptr) v = y[static_cast<size_t>(r) * ldy + c] > 0.f ? v * scale : 0.f;
    }
    dz[static_cast<size_t>(r) * lddz + c] = v;
  }
}

// Batch assembly out of a DEVICE-RESIDENT uint8 dataset: out[r, :] = src[idx[r], :] (row_bytes a multiple of 16: one 16-byte
// vector per thread), and y_out[r] = labels[idx[r]] widened to int64. The whole FashionMNIST training set is 47 MB, a rounding
// error of 180 GB of HBM, so the per-epoch shuffle costs one index upload and this kernel at HBM speed instead of a host-side
// gather of 784-byte rows.
__global__ void __launch_bounds__(256)
u8_gather_rows_kernel(const uint8_t* __restrict__ src, const int32_t* __restrict__ idx, int rows, int vecs_per_row,
                      uint8_t* __restrict__ out, const uint8_t* __restrict__ labels, long long* __restrict__ y_out) {
  const size_t total = static_cast<size_t>(rows) * vecs_per_row;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < total; i += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int r = static_cast<int>(i / vecs_per_row), v = static_cast<int>(i - static_cast<size_t>(r) * vecs_per_row);
    const size_t s = static_cast<size_t>(__ldg(idx + r));
    reinterpret_cast<uint4*>(out)[i] = __ldg(reinterpret_cast<const uint4*>(src) + s * vecs_per_row + v);
    if (v == 0 && labels != nullptr) y_out[r] = static_cast<long long>(labels[s]);
  }
}

// Adam / AdamW for bf16 parameters: the update runs on an fp32 master copy (and fp32 moments); the bf16 parameter the model
// computes with is re-derived from the master every step, the bf16 gradient comes straight from the (all-reduced) gradient arena.
// One pass over 2 + 2 + 3 x 4 bytes per element, eight elements (one 16-byte bf16 vector) per thread.
__global__ void __launch_bounds__(256)
adam_bf16_kernel(__nv_bfloat16* __restrict__ p16, const __nv_bfloat16* __restrict__ g16, float* __restrict__ master, float* __restrict__ m,
                 float* __restrict__ v, size_t n, const rtdc_adam_hyper* __restrict__ hp) {
  const rtdc_adam_hyper h = *hp;
  const float t = static_cast<float>(h.step);
  const float bc1 = 1.f - powf(h.beta1, t);
  const float bc2_sqrt = sqrtf(1.f - powf(h.beta2, t));
  const size_t nvec = n / 8;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; i < nvec; i += stride) {
    const uint4 graw = __ldg(reinterpret_cast<const uint4*>(g16) + i);
    const uint32_t gw[4] = {graw.x, graw.y, graw.z, graw.w};
    float pv[8], mv[8], vv[8];
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      const float4 a = reinterpret_cast<const float4*>(master)[2 * i + q];
      const float4 b = reinterpret_cast<const float4*>(m)[2 * i + q];
      const float4 c = reinterpret_cast<const float4*>(v)[2 * i + q];
      pv[4 * q] = a.x; pv[4 * q + 1] = a.y; pv[4 * q + 2] = a.z; pv[4 * q + 3] = a.w;
      mv[4 * q] = b.x; mv[4 * q + 1] = b.y; mv[4 * q + 2] = b.z; mv[4 * q + 3] = b.w;
      vv[4 * q] = c.x; vv[4 * q + 1] = c.y; vv[4 * q + 2] = c.z; vv[4 * q + 3] = c.w;
    }
    uint32_t out[4];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      adam_update(pv[2 * q], __uint_as_float(gw[q] << 16), mv[2 * q], vv[2 * q], h, bc1, bc2_sqrt);
      adam_update(pv[2 * q + 1], __uint_as_float(gw[q] & 0xFFFF0000u), mv[2 * q + 1], vv[2 * q + 1], h, bc1, bc2_sqrt);
      const __nv_bfloat162 r = __floats2bfloat162_rn(pv[2 * q], pv[2 * q + 1]);
      out[q] = *reinterpret_cast<const uint32_t*>(&r);
    }
#pragma unroll
    for (int q = 0; q < 2; ++q) {
      reinterpret_cast<float4*>(master)[2 * i + q] = make_float4(pv[4 * q], pv[4 * q + 1], pv[4 * q + 2], pv[4 * q + 3]);
      reinterpret_cast<float4*>(m)[2 * i + q] = make_float4(mv[4 * q], mv[4 * q + 1], mv[4 * q + 2], mv[4 * q + 3]);
      reinterpret_cast<float4*>(v)[2 * i + q] = make_float4(vv[4 * q], vv[4 * q + 1], vv[4 * q + 2], vv[4 * q + 3]);
    }
    reinterpret_cast<uint4*>(p16)[i] = make_uint4(out[0], out[1], out[2], out[3]);
  }
  if (blockIdx.x == 0) {
    for (size_t i = nvec * 8 + threadIdx.x; i < n; i += blockDim.x) {
      float pf = master[i], mf = m[i], vf = v[i];
      adam_update(pf, __bfloat162float(g16[i]), mf, vf, h, bc1, bc2_sqrt);
      master[i] = pf; m[i] = mf; v[i] = vf;
      p16[i] = __float2bfloat16_rn(pf);
    }
  }
}

__global__ void adam_advance_kernel(rtdc_adam_hyper* h) { h->step += 1; }
__global__ void counter_advance_kernel(uint32_t* c, rtdc_sgd_hyper* h) {
  if (c) c[0] += 1u;
  if (h) h->first_step = 0;
}

static int grid_for(size_t work_items, int threads, int per_sm) {
  size_t blocks = (work_items + threads - 1) / threads;
  const size_t cap = static_cast<size_t>(148) * per_sm;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return static_cast<int>(blocks);
}

}  // namespace rtdc

using namespace rtdc;

RTDC_API const char* rtdc_last_error(void) { return error_buffer(); }
RTDC_API int rtdc_version(void) { return 1; }
RTDC_API const char* rtdc_arch(void) { return "sm_100a"; }

RTDC_API size_t rtdc_softmax_ce_scratch_bytes(int) { return sizeof(CeScratch); }

RTDC_API int rtdc_softmax_ce(const float* logits, int ldz, const int64_t* labels, int rows, int classes,
                             float* loss_out, float* dlogits, float grad_scale, int relu_mask,
                             unsigned long long* correct_accum, float* loss_sum_accum, float* dbias_out, void* scratch,
                             size_t scratch_bytes, void* stream) {
  if (!logits || !labels || rows <= 0) return fail(RTDC_ERR_ARG, "logits/labels null or rows <= 0");
  if (classes < 1 || classes > 32 || ldz < classes) return fail(RTDC_ERR_LIMIT, "softmax_ce supports 1..32 classes with ldz >= classes (got %d, ldz %d)", classes, ldz);
  if ((ldz & 3) == 0 && ldz > 32) return fail(RTDC_ERR_LIMIT, "softmax_ce: ldz must be <= 32 (got %d)", ldz);
  if (!scratch || scratch_bytes < sizeof(CeScratch)) return fail(RTDC_ERR_ARG, "softmax_ce needs %zu bytes of scratch", sizeof(CeScratch));
  const int blocks = grid_for(static_cast<size_t>(rows), CE_THREADS, 8);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (classes <= 16 && ldz <= 16)
    softmax_ce_kernel<16><<<blocks, CE_THREADS, 0, st>>>(logits, ldz, reinterpret_cast<const long long*>(labels), rows, classes, loss_out, dlogits, grad_scale, relu_mask, correct_accum, loss_sum_accum, dbias_out, static_cast<CeScratch*>(scratch));
  else
    softmax_ce_kernel<32><<<blocks, CE_THREADS, 0, st>>>(logits, ldz, reinterpret_cast<const long long*>(labels), rows, classes, loss_out, dlogits, grad_scale, relu_mask, correct_accum, loss_sum_accum, dbias_out, static_cast<CeScratch*>(scratch));
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

static void colsum_shape(int rows, int cols, int* col_groups, int* slabs, int* rows_per_slab) {
  *col_groups = ceil_div(cols, CS_COLS);
  int s = (148 * 4) / *col_groups;
  if (s > CS_MAX_SLABS) s = CS_MAX_SLABS;
  const int max_by_rows = ceil_div(rows, 64);
  if (s > max_by_rows) s = max_by_rows;
  if (s < 1) s = 1;
  *rows_per_slab = ceil_div(rows, s);
  *slabs = ceil_div(rows, *rows_per_slab);
}

constexpr size_t CS_COUNTER_BYTES = 4096;  // counters first, at a fixed place (see gemm_tf32.cu)
RTDC_API size_t rtdc_colsum_scratch_bytes(int rows, int cols) {
  int cg, slabs, rps;
  colsum_shape(rows, cols, &cg, &slabs, &rps);
  return CS_COUNTER_BYTES + static_cast<size_t>(slabs) * cg * CS_COLS * 4;
}

RTDC_API int rtdc_colsum(const float* x, int ld, int rows, int cols, float* out, void* scratch, size_t scratch_bytes,
                         void* stream) {
  if (!x || !out || rows <= 0 || cols <= 0 || ld < cols) return fail(RTDC_ERR_ARG, "colsum: bad arguments");
  int cg, slabs, rps;
  colsum_shape(rows, cols, &cg, &slabs, &rps);
  const size_t need = rtdc_colsum_scratch_bytes(rows, cols);
  if (!scratch || scratch_bytes < need) return fail(RTDC_ERR_ARG, "colsum needs %zu bytes of scratch (got %zu)", need, scratch_bytes);
  if (static_cast<size_t>(cg) * 4 > CS_COUNTER_BYTES) return fail(RTDC_ERR_LIMIT, "colsum supports at most %zu columns", CS_COUNTER_BYTES / 4 * CS_COLS);
  unsigned int* counters = static_cast<unsigned int*>(scratch);
  float* partial = reinterpret_cast<float*>(static_cast<uint8_t*>(scratch) + CS_COUNTER_BYTES);
  RTDC_CUDA_CHECK(launch_pdl(colsum_kernel, dim3(cg, slabs), dim3(CS_THREADS), 0, static_cast<cudaStream_t>(stream), x, ld, rows, cols, out, partial,
                             counters, rps));
  return RTDC_OK;
}

// ---------------------------------------------------------------------------------------------
// Column sums for callers that must not allocate device memory.
//
// Who this is for: hosts that handed out all of their device memory up front and keep it that way —
// an inference server whose arena already owns every free byte of HBM (a cudaMalloc there simply
// fails), a real-time loop that cannot afford the device-wide synchronisation cudaMalloc/cudaFree
// imply, or a plain C/C++ embedder that has no caching allocator to borrow a workspace from. The
// regular rtdc_colsum makes such a caller find `rtdc_colsum_scratch_bytes` of zeroed device memory;
// this variant brings its own: the partial sums and arrival counters live in a file-scope __device__
// array that is part of the library image (each GPU that loads the library carries it until exit).
constexpr int COLSUM_NOALLOC_MAX_COLS = RTDC_COLSUM_NOALLOC_MAX_COLS;
constexpr size_t COLSUM_NOALLOC_SCRATCH_BYTES =
    CS_COUNTER_BYTES + static_cast<size_t>(CS_MAX_SLABS) * (COLSUM_NOALLOC_MAX_COLS / CS_COLS) * CS_COLS * sizeof(float);
static_assert(COLSUM_NOALLOC_SCRATCH_BYTES == RTDC_COLSUM_NOALLOC_SCRATCH_BYTES, "header and kernel disagree on the built-in scratch size");
static __device__ __align__(256) unsigned char g_colsum_noalloc_scratch[COLSUM_NOALLOC_SCRATCH_BYTES];

constexpr int NOALLOC_MAX_DEVICES = 64;
static void* g_colsum_noalloc_ptr[NOALLOC_MAX_DEVICES] = {nullptr};

// In host code `g_colsum_noalloc_scratch` is only a handle for the symbol, not a device address; the
// address has to be asked for, once per device, with that device current.
static int colsum_noalloc_scratch(void** out) {
  int dev = -1;
  RTDC_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= NOALLOC_MAX_DEVICES) return rtdc::fail(rtdc::RTDC_ERR_LIMIT, "colsum_noalloc: device ordinal %d not supported (max %d)", dev, NOALLOC_MAX_DEVICES - 1);
  void* p = __atomic_load_n(&g_colsum_noalloc_ptr[dev], __ATOMIC_ACQUIRE);
  if (p == nullptr) {
    RTDC_CUDA_CHECK(cudaGetSymbolAddress(&p, g_colsum_noalloc_scratch));
    if (p == nullptr) return rtdc::fail(rtdc::RTDC_ERR_CUDA, "colsum_noalloc: cudaGetSymbolAddress returned a null address");
    __atomic_store_n(&g_colsum_noalloc_ptr[dev], p, __ATOMIC_RELEASE);  // racing first calls resolve the same address
  }
  *out = p;
  return rtdc::RTDC_OK;
}

RTDC_API int rtdc_colsum_noalloc(const float* x, int ld, int rows, int cols, float* out, void* stream) {
  if (!x || !out || rows <= 0 || cols <= 0 || ld < cols) return fail(RTDC_ERR_ARG, "colsum_noalloc: bad arguments");
  if (cols > COLSUM_NOALLOC_MAX_COLS)
    return fail(RTDC_ERR_LIMIT, "colsum_noalloc: %d columns exceed the built-in scratch (at most %d columns; use rtdc_colsum with "
                "caller-provided scratch for wider matrices)", cols, COLSUM_NOALLOC_MAX_COLS);
  int cg, slabs, rps;
  colsum_shape(rows, cols, &cg, &slabs, &rps);
  const size_t need = CS_COUNTER_BYTES + static_cast<size_t>(slabs) * cg * CS_COLS * sizeof(float);
  if (need > COLSUM_NOALLOC_SCRATCH_BYTES)
    return fail(RTDC_ERR_LIMIT, "colsum_noalloc: needs %zu bytes of scratch, built-in scratch holds %zu", need, COLSUM_NOALLOC_SCRATCH_BYTES);
  void* scratch = nullptr;
  int rc = colsum_noalloc_scratch(&scratch);
  if (rc != RTDC_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // The array is shared by every call and still holds what the previous one left. The kernel only needs
  // its arrival counters to start at zero (it overwrites the partial sums before reading them); clear
  // them here rather than trusting that the last call ran to completion.
  RTDC_CUDA_CHECK(cudaMemsetAsync(scratch, 0, CS_COUNTER_BYTES, st));
  unsigned int* counters = static_cast<unsigned int*>(scratch);
  float* partial = reinterpret_cast<float*>(static_cast<unsigned char*>(scratch) + CS_COUNTER_BYTES);
  colsum_kernel<<<dim3(cg, slabs), CS_THREADS, 0, st>>>(x, ld, rows, cols, out, partial, counters, rps);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API int rtdc_u8_normalize(const uint8_t* src, float* dst, size_t n, float mean, float std, void* stream) {
  if (!src || !dst) return fail(RTDC_ERR_ARG, "u8_normalize: null pointer");
  if (n == 0) return RTDC_OK;
  if ((reinterpret_cast<uintptr_t>(src) & 15u) || (reinterpret_cast<uintptr_t>(dst) & 15u)) return fail(RTDC_ERR_ARG, "u8_normalize: pointers must be 16-byte aligned");
  const float scale = 1.0f / (255.0f * std), shift = -mean / std;
  u8_normalize_kernel<<<grid_for(n / 16 + 1, 256, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(src, dst, n, scale, shift);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API int rtdc_u8_affine(const uint8_t* src, float* dst, size_t n, float scale, float shift, void* stream) {
  if (!src || !dst) return fail(RTDC_ERR_ARG, "u8_affine: null pointer");
  if (n == 0) return RTDC_OK;
  if ((reinterpret_cast<uintptr_t>(src) & 15u) || (reinterpret_cast<uintptr_t>(dst) & 15u)) return fail(RTDC_ERR_ARG, "u8_affine: pointers must be 16-byte aligned");
  u8_normalize_kernel<<<grid_for(n / 16 + 1, 256, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(src, dst, n, scale, shift);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API int rtdc_u8_gather_rows(const uint8_t* src, const int32_t* idx, int rows, int row_bytes, uint8_t* out,
                                 const uint8_t* labels_or_null, int64_t* y_out_or_null, void* stream) {
  if (!src || !idx || !out) return fail(RTDC_ERR_ARG, "u8_gather_rows: null pointer");
  if (rows <= 0) return RTDC_OK;
  if (row_bytes <= 0 || row_bytes % 16 != 0) return fail(RTDC_ERR_ARG, "u8_gather_rows: row_bytes must be a positive multiple of 16 (got %d)", row_bytes);
  if ((reinterpret_cast<uintptr_t>(src) & 15u) || (reinterpret_cast<uintptr_t>(out) & 15u)) return fail(RTDC_ERR_ARG, "u8_gather_rows: pointers must be 16-byte aligned");
  if (labels_or_null && !y_out_or_null) return fail(RTDC_ERR_ARG, "u8_gather_rows: labels without an output");
  const int vpr = row_bytes / 16;
  u8_gather_rows_kernel<<<grid_for(static_cast<size_t>(rows) * vpr, 256, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      src, idx, rows, vpr, out, labels_or_null, reinterpret_cast<long long*>(y_out_or_null));
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

// Batch assembly by the copy engines: n pieces of a pinned host dataset (byte offsets src_offsets[i], sizes[i] bytes each) land
// back to back in dst_dev. One cudaMemcpyAsync per piece, issued from native code so that a step's ~40 copies cost the host
// tens of microseconds, not a Python loop.
RTDC_API int rtdc_h2d_pieces(void* dst_dev, const void* src_host_base, const int64_t* src_offsets, const int64_t* sizes, int n,
                             void* stream) {
  if (!dst_dev || !src_host_base || ((!src_offsets || !sizes) && n > 0)) return fail(RTDC_ERR_ARG, "h2d_pieces: null pointer");
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  char* dst = static_cast<char*>(dst_dev);
  const char* src = static_cast<const char*>(src_host_base);
  for (int i = 0; i < n; ++i) {
    if (sizes[i] <= 0) continue;
    RTDC_CUDA_CHECK(cudaMemcpyAsync(dst, src + src_offsets[i], static_cast<size_t>(sizes[i]), cudaMemcpyHostToDevice, st));
    dst += sizes[i];
  }
  return RTDC_OK;
}

RTDC_API int rtdc_act_backward(const float* dy, int lddy, const float* y_or_null, int ldy, int rows, int cols, int cols_out, float scale,
                               float* dz, int lddz, void* stream) {
  if (!dy || !dz) return fail(RTDC_ERR_ARG, "act_backward: null pointer");
  if (rows <= 0 || cols <= 0 || cols_out < cols || lddz < cols_out || lddy < cols || (y_or_null && ldy < cols))
    return fail(RTDC_ERR_ARG, "act_backward: bad shape (rows %d cols %d cols_out %d)", rows, cols, cols_out);
  act_backward_kernel<<<grid_for(static_cast<size_t>(rows) * cols_out, 256, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      dy, lddy, y_or_null, ldy, rows, cols, cols_out, scale, dz, lddz);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API int rtdc_u8_affine_paced(const uint8_t* src, float* dst, size_t n, float scale, float shift, int blocks, void* stream) {
  if (!src || !dst) return fail(RTDC_ERR_ARG, "u8_affine: null pointer");
  if (n == 0) return RTDC_OK;
  if ((reinterpret_cast<uintptr_t>(src) & 15u) || (reinterpret_cast<uintptr_t>(dst) & 15u)) return fail(RTDC_ERR_ARG, "u8_affine: pointers must be 16-byte aligned");
  int grid = grid_for(n / 16 + 1, 256, 8);
  if (blocks > 0 && blocks < grid) grid = blocks;
  u8_normalize_kernel<<<grid, 256, 0, static_cast<cudaStream_t>(stream)>>>(src, dst, n, scale, shift);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

namespace rtdc {
// out[n] = bias[n] + t * sum_k w[n, k]: one warp per row, 4 independent float4 loads per lane in flight,
// fixed summation order (lane-strided partials, then a shuffle tree).
__global__ void __launch_bounds__(256)
fold_input_bias_kernel(const float* __restrict__ w, int ldw, int rows, int cols, const float* __restrict__ bias, float t,
                       float* __restrict__ out) {
  const int row = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (row >= rows) return;
  const float* wr = w + static_cast<size_t>(row) * ldw;
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  if ((ldw & 3) == 0 && (reinterpret_cast<uintptr_t>(w) & 15u) == 0) {
    const int nv = cols >> 2;
    for (int i0 = lane; i0 < nv; i0 += 128) {
      float4 v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u)
        v[u] = (i0 + 32 * u < nv) ? __ldg(reinterpret_cast<const float4*>(wr) + i0 + 32 * u) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
      for (int u = 0; u < 4; ++u) acc[u] += (v[u].x + v[u].y) + (v[u].z + v[u].w);
    }
    for (int k = (nv << 2) + lane; k < cols; k += 32) acc[0] += __ldg(wr + k);
  } else {
    for (int k = lane; k < cols; k += 32) acc[0] += __ldg(wr + k);
  }
  float sum = (acc[0] + acc[1]) + (acc[2] + acc[3]);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
  if (lane == 0) out[row] = fmaf(t, sum, bias != nullptr ? __ldg(bias + row) : 0.f);
}
}  // namespace rtdc

RTDC_API int rtdc_fold_input_bias(const float* w, int ldw, int rows, int cols, const float* bias, float t, float* out, void* stream) {
  if (!w || !out) return fail(RTDC_ERR_ARG, "fold_input_bias: null pointer");
  if (rows <= 0 || cols <= 0 || ldw < cols) return fail(RTDC_ERR_ARG, "fold_input_bias: bad shape (%d x %d, ld %d)", rows, cols, ldw);
  rtdc::fold_input_bias_kernel<<<(rows + 7) / 8, 256, 0, static_cast<cudaStream_t>(stream)>>>(w, ldw, rows, cols, bias, t, out);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

static int sgd_launch(float* p, const float* g, float* m, size_t n, rtdc_sgd_hyper* hyper_dev, uint32_t* counter, int advance, void* stream,
                      const float* loss = nullptr, float* loss_log = nullptr, uint32_t log_slots = 0) {
  if (log_slots & (log_slots - 1)) return fail(RTDC_ERR_ARG, "sgd_step: loss_log_slots must be a power of two");
  if (!p || !g || !hyper_dev) return fail(RTDC_ERR_ARG, "sgd_step: null pointer");
  if (n == 0) return RTDC_OK;
  if ((reinterpret_cast<uintptr_t>(p) & 15u) || (reinterpret_cast<uintptr_t>(g) & 15u) || (reinterpret_cast<uintptr_t>(m) & 15u))
    return fail(RTDC_ERR_ARG, "sgd_step: arenas must be 16-byte aligned");
  RTDC_CUDA_CHECK(launch_pdl(sgd_kernel, dim3(grid_for(n / 4 + 1, 256, 8)), dim3(256), 0, static_cast<cudaStream_t>(stream), p, g, m, n, hyper_dev,
                             counter, advance, loss, log_slots ? loss_log : nullptr, log_slots ? log_slots - 1 : 0u));
  return RTDC_OK;
}

RTDC_API int rtdc_sgd_step(float* p, const float* g, float* m, size_t n, const rtdc_sgd_hyper* hyper_dev, void* stream) {
  return sgd_launch(p, g, m, n, const_cast<rtdc_sgd_hyper*>(hyper_dev), nullptr, 0, stream);
}

RTDC_API int rtdc_sgd_step_advance(float* p, const float* g, float* m, size_t n, rtdc_sgd_hyper* hyper_dev, uint32_t* step_counter_or_null,
                                   void* stream) {
  return sgd_launch(p, g, m, n, hyper_dev, step_counter_or_null, 1, stream);
}

RTDC_API int rtdc_sgd_step_advance_log(float* p, const float* g, float* m, size_t n, rtdc_sgd_hyper* hyper_dev, uint32_t* step_counter_or_null,
                                       const float* loss_dev_or_null, float* loss_log_host_or_null, uint32_t loss_log_slots, void* stream) {
  return sgd_launch(p, g, m, n, hyper_dev, step_counter_or_null, 1, stream, loss_dev_or_null, loss_log_host_or_null, loss_log_slots);
}

RTDC_API int rtdc_adam_step(float* p, const float* g, float* m, float* v, size_t n, const rtdc_adam_hyper* hyper_dev, void* stream) {
  if (!p || !g || !m || !v || !hyper_dev) return fail(RTDC_ERR_ARG, "adam_step: null pointer");
  if (n == 0) return RTDC_OK;
  if ((reinterpret_cast<uintptr_t>(p) & 15u) || (reinterpret_cast<uintptr_t>(g) & 15u) || (reinterpret_cast<uintptr_t>(m) & 15u) || (reinterpret_cast<uintptr_t>(v) & 15u))
    return fail(RTDC_ERR_ARG, "adam_step: arenas must be 16-byte aligned");
  adam_kernel<<<grid_for(n / 4 + 1, 256, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(p, g, m, v, n, hyper_dev);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API int rtdc_adam_step_bf16(void* p_bf16, const void* g_bf16, float* master, float* m, float* v, size_t n,
                                 const rtdc_adam_hyper* hyper_dev, void* stream) {
  if (!p_bf16 || !g_bf16 || !master || !m || !v || !hyper_dev) return fail(RTDC_ERR_ARG, "adam_step_bf16: null pointer");
  if (n == 0) return RTDC_OK;
  const uintptr_t all = reinterpret_cast<uintptr_t>(p_bf16) | reinterpret_cast<uintptr_t>(g_bf16) | reinterpret_cast<uintptr_t>(master) |
                        reinterpret_cast<uintptr_t>(m) | reinterpret_cast<uintptr_t>(v);
  if (all & 15u) return fail(RTDC_ERR_ARG, "adam_step_bf16: arenas must be 16-byte aligned");
  adam_bf16_kernel<<<grid_for(n / 8 + 1, 256, 8), 256, 0, static_cast<cudaStream_t>(stream)>>>(
      static_cast<__nv_bfloat16*>(p_bf16), static_cast<const __nv_bfloat16*>(g_bf16), master, m, v, n, hyper_dev);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API int rtdc_adam_advance(rtdc_adam_hyper* hyper_dev, void* stream) {
  adam_advance_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(hyper_dev);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API int rtdc_counter_advance(uint32_t* counter_dev, rtdc_sgd_hyper* sgd_hyper_or_null, void* stream) {
  counter_advance_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream)>>>(counter_dev, sgd_hyper_or_null);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}
