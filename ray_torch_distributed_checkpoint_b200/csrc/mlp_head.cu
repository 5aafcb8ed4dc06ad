// Fused "head" of the reference MLP for training: the last Linear (hidden -> classes, classes <= 16),
// its trailing ReLU, the softmax cross-entropy loss and ALL of that layer's backward in one pass over
// the hidden activations (K3 + K4 + K5 + the layer-3 part of K6 in SURVEY.md §2.5; reference:
// my_ray_module.py:105-106 forward, :156 loss, :159 backward).
//
// Why a dedicated kernel: with 10 output classes these ops are 1.3 % of the step's FLOPs but, run as
// three skinny tensor-core GEMMs + a loss kernel + reductions, they were 16 % of its time — each of them
// re-reads the 39 MB activation matrix and pays a full pipeline prologue for almost no math. Here a
// block loads a 32-row tile of H once into shared memory and produces, in exact fp32 FFMA arithmetic:
//   z = relu(H W^T + b)            (never written out)
//   loss_i, dz = (softmax(z) - onehot) * grad_scale * 1[z > 0]
//   dH = (dz W) * 1[H > 0] * mask_scale      -> global (the only large write)
//   dW += dz^T H,  db += sum_rows dz          -> per-block partials, summed in block order by head_reduce
#include "common.h"
#include "ptx.cuh"
#include "../../include/rtdc_b200.h"

namespace rtdc {

// Two blocks per SM (89 KB of smem each at K = 512): while one block waits for its H tile to land, the other computes,
// so the tile load needs no double buffering inside a block.
constexpr int HD_ROWS = 32;       // rows of H per tile
constexpr int HD_THREADS = 256;     // 8 warps per block, 2 blocks per SM
constexpr int HD_GROUPS = HD_THREADS / 128;   // thread groups of 128 (one float4 column slice per thread) in phases C/D
constexpr int HD_BLOCKS_PER_SM = 2;
constexpr int HD_MAXK = 512;      // hidden width limit (smem tile 32 x 512 fp32 = 64 KB)
constexpr int HD_PAD = 4;         // row padding (floats) -> conflict-free 128-bit accesses across rows
constexpr int HD_DZ_LD = 16;      // dz rows padded to 16 floats

struct HeadParams {
  const float* h; int ldh;
  const float* w;                 // [C, K] row-major
  const float* bias;              // [C]
  const long long* labels;
  int rows, K;
  float grad_scale, mask_scale;
  float* dh; int lddh;
  float* dw_partial;              // [grid][C][K]
  float* db_partial;              // [grid][16]
  float* loss_partial;            // [grid]
  float* dbh_partial;             // [grid][K]: column sums of dh (bias gradient of the layer below), or null
};

template <int C>
__global__ void __launch_bounds__(HD_THREADS, HD_BLOCKS_PER_SM)
mlp_head_train_kernel(const HeadParams p) {
  pdl_launch_dependents();
  pdl_wait();
  extern __shared__ __align__(16) float smem[];
  const int K = p.K, K4 = K >> 2, LD = K + HD_PAD;
  float* sh = smem;                        // [HD_ROWS][LD]
  float* sw = sh + HD_ROWS * LD;           // [C][LD]
  float* sdz = sw + C * LD;                // [HD_ROWS][HD_DZ_LD]
  float* sb = sdz + HD_ROWS * HD_DZ_LD;    // [16]
  __shared__ float s_red[HD_THREADS / 32];

  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  // weights + bias -> smem once per block
  for (int i = t; i < C * K4; i += HD_THREADS) {
    const int c = i / K4, k4 = i - c * K4;
    *reinterpret_cast<float4*>(sw + c * LD + 4 * k4) = __ldg(reinterpret_cast<const float4*>(p.w + static_cast<size_t>(c) * K) + k4);
  }
  if (t < 16) sb[t] = (t < C && p.bias != nullptr) ? __ldg(p.bias + t) : 0.f;

  // persistent accumulators
  constexpr int CH = (C + HD_GROUPS - 1) / HD_GROUPS;   // classes per thread group in the dW phase
  float4 dw_acc[CH];
#pragma unroll
  for (int c = 0; c < CH; ++c) dw_acc[c] = make_float4(0.f, 0.f, 0.f, 0.f);
  float db_acc = 0.f;                      // thread t < C owns class t
  float4 dbh_acc = make_float4(0.f, 0.f, 0.f, 0.f);   // column sums of dh over this thread's rows
  float loss_acc = 0.f;

  const int tiles = (p.rows + HD_ROWS - 1) / HD_ROWS;
  for (int tile = blockIdx.x; tile < tiles; tile += gridDim.x) {
    const int row0 = tile * HD_ROWS;
    __syncthreads();                       // previous tile's readers are done with sh / sdz
    // ---- A: H tile -> smem with cp.async (16 bytes each, every load of the tile in flight at once)
    {
      const int total = HD_ROWS * K4;
      for (int i = t; i < total; i += HD_THREADS) {
        const int r = i / K4, k4 = i - r * K4;
        float* dst = sh + r * LD + 4 * k4;
        if (row0 + r < p.rows) {
          const float* src = p.h + static_cast<size_t>(row0 + r) * p.ldh + 4 * k4;
          asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(static_cast<uint32_t>(__cvta_generic_to_shared(dst))), "l"(src) : "memory");
        } else {
          *reinterpret_cast<float4*>(dst) = make_float4(0.f, 0.f, 0.f, 0.f);
        }
      }
      asm volatile("cp.async.commit_group;" ::: "memory");
      asm volatile("cp.async.wait_group 0;" ::: "memory");
    }
    __syncthreads();

    // ---- B: logits for two rows per thread (16 threads share a row pair and stride over K), then CE
    {
      const int g = t & 15, rp = t >> 4;   // rp in 0..HD_ROWS/2-1: rows rp and rp + HD_ROWS/2
      static_assert(HD_THREADS / 16 == HD_ROWS / 2, "phase B maps 16 threads to a pair of rows");
      float a0[C], a1[C];
#pragma unroll
      for (int c = 0; c < C; ++c) { a0[c] = 0.f; a1[c] = 0.f; }
      const float* h0 = sh + rp * LD;
      const float* h1 = sh + (rp + HD_ROWS / 2) * LD;
      for (int k4 = g; k4 < K4; k4 += 16) {
        const float4 x0 = *reinterpret_cast<const float4*>(h0 + 4 * k4);
        const float4 x1 = *reinterpret_cast<const float4*>(h1 + 4 * k4);
#pragma unroll
        for (int c = 0; c < C; ++c) {
          const float4 wv = *reinterpret_cast<const float4*>(sw + c * LD + 4 * k4);
          a0[c] = fmaf(x0.x, wv.x, fmaf(x0.y, wv.y, fmaf(x0.z, wv.z, fmaf(x0.w, wv.w, a0[c]))));
          a1[c] = fmaf(x1.x, wv.x, fmaf(x1.y, wv.y, fmaf(x1.z, wv.z, fmaf(x1.w, wv.w, a1[c]))));
        }
      }
#pragma unroll
      for (int c = 0; c < C; ++c) {
#pragma unroll
        for (int o = 1; o < 16; o <<= 1) {
          a0[c] += __shfl_xor_sync(0xffffffffu, a0[c], o);
          a1[c] += __shfl_xor_sync(0xffffffffu, a1[c], o);
        }
      }
      if (g < 2) {                         // g == 0 finishes row rp, g == 1 finishes row rp + HD_ROWS/2
        const int r = rp + (HD_ROWS / 2) * g;
        const int row = row0 + r;
        float z[C];
        float mx = -INFINITY;
#pragma unroll
        for (int c = 0; c < C; ++c) {
          z[c] = fmaxf((g == 0 ? a0[c] : a1[c]) + sb[c], 0.f);
          mx = fmaxf(mx, z[c]);
        }
        float sum = 0.f, e[C];
#pragma unroll
        for (int c = 0; c < C; ++c) { e[c] = __expf(z[c] - mx); sum += e[c]; }
        float* dzr = sdz + r * HD_DZ_LD;
        if (row < p.rows) {
          const long long y = p.labels[row];
          float zy = 0.f;
          const float inv = p.grad_scale / sum;
#pragma unroll
          for (int c = 0; c < C; ++c) {
            if (c == y) zy = z[c];
            const float d = e[c] * inv - ((c == y) ? p.grad_scale : 0.f);
            dzr[c] = (z[c] > 0.f) ? d : 0.f;
          }
          loss_acc += (mx + __logf(sum)) - zy;
        } else {
#pragma unroll
          for (int c = 0; c < C; ++c) dzr[c] = 0.f;
        }
#pragma unroll
        for (int c = C; c < HD_DZ_LD; ++c) dzr[c] = 0.f;
      }
    }
    __syncthreads();

    // ---- db: class t sums its column of dz
    if (t < C) {
      float sacc = 0.f;
      for (int r = 0; r < HD_ROWS; ++r) sacc += sdz[r * HD_DZ_LD + t];
      db_acc += sacc;
    }

    const int k4 = t & 127, grp = t >> 7;  // groups of 128 threads; thread owns columns 4*k4..4*k4+3
    constexpr int RPG = HD_ROWS / HD_GROUPS;   // rows per group in phase C
    if (k4 < K4) {
      // ---- C: dH = (dz W) masked by H > 0, rows grp*RPG .. grp*RPG+RPG-1
      {
        float4 wr[C];
#pragma unroll
        for (int c = 0; c < C; ++c) wr[c] = *reinterpret_cast<const float4*>(sw + c * LD + 4 * k4);
        for (int r = grp * RPG; r < grp * RPG + RPG; ++r) {
          const int row = row0 + r;
          if (row >= p.rows) break;
          float d[HD_DZ_LD];
#pragma unroll
          for (int q = 0; q < (C + 3) / 4; ++q) {
            const float4 tq = *reinterpret_cast<const float4*>(sdz + r * HD_DZ_LD + 4 * q);
            d[4 * q] = tq.x; d[4 * q + 1] = tq.y; d[4 * q + 2] = tq.z; d[4 * q + 3] = tq.w;
          }
          const float4 hv = *reinterpret_cast<const float4*>(sh + r * LD + 4 * k4);
          float4 o = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int c = 0; c < C; ++c) {
            o.x = fmaf(d[c], wr[c].x, o.x); o.y = fmaf(d[c], wr[c].y, o.y);
            o.z = fmaf(d[c], wr[c].z, o.z); o.w = fmaf(d[c], wr[c].w, o.w);
          }
          o.x = hv.x > 0.f ? o.x * p.mask_scale : 0.f;
          o.y = hv.y > 0.f ? o.y * p.mask_scale : 0.f;
          o.z = hv.z > 0.f ? o.z * p.mask_scale : 0.f;
          o.w = hv.w > 0.f ? o.w * p.mask_scale : 0.f;
          *reinterpret_cast<float4*>(p.dh + static_cast<size_t>(row) * p.lddh + 4 * k4) = o;
          dbh_acc.x += o.x; dbh_acc.y += o.y; dbh_acc.z += o.z; dbh_acc.w += o.w;
        }
      }
      // ---- D: dW[c][k] += sum_r dz[r][c] * H[r][k], classes grp*CH .. grp*CH+CH-1
      for (int r = 0; r < HD_ROWS; ++r) {
        const float4 hv = *reinterpret_cast<const float4*>(sh + r * LD + 4 * k4);
#pragma unroll
        for (int c = 0; c < CH; ++c) {
          const int cc = grp * CH + c;
          const float d = cc < C ? sdz[r * HD_DZ_LD + cc] : 0.f;
          dw_acc[c].x = fmaf(d, hv.x, dw_acc[c].x); dw_acc[c].y = fmaf(d, hv.y, dw_acc[c].y);
          dw_acc[c].z = fmaf(d, hv.z, dw_acc[c].z); dw_acc[c].w = fmaf(d, hv.w, dw_acc[c].w);
        }
      }
    }
  }

  // ---- block partials
  {
    const int k4 = t & 127, grp = t >> 7;
    if (k4 < K4) {
#pragma unroll
      for (int c = 0; c < CH; ++c) {
        const int cc = grp * CH + c;
        if (cc < C)
          *reinterpret_cast<float4*>(p.dw_partial + (static_cast<size_t>(blockIdx.x) * C + cc) * K + 4 * k4) = dw_acc[c];
      }
    }
    if (p.dbh_partial != nullptr) {
      __syncthreads();                     // everyone is done with the H tile: reuse it as a [groups][K] staging area
      if (k4 < K4) *reinterpret_cast<float4*>(sh + grp * LD + 4 * k4) = dbh_acc;
      __syncthreads();
      if (grp == 0 && k4 < K4) {
        float4 tot = *reinterpret_cast<const float4*>(sh + 4 * k4);
#pragma unroll
        for (int g2 = 1; g2 < HD_GROUPS; ++g2) {
          const float4 u = *reinterpret_cast<const float4*>(sh + g2 * LD + 4 * k4);
          tot.x += u.x; tot.y += u.y; tot.z += u.z; tot.w += u.w;
        }
        *reinterpret_cast<float4*>(p.dbh_partial + static_cast<size_t>(blockIdx.x) * K + 4 * k4) = tot;
      }
    }
    if (t < 16) p.db_partial[blockIdx.x * 16 + t] = t < C ? db_acc : 0.f;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) loss_acc += __shfl_xor_sync(0xffffffffu, loss_acc, o);
    if (lane == 0) s_red[warp] = loss_acc;
    __syncthreads();
    if (t == 0) {
      float tot = 0.f;
      for (int wv = 0; wv < HD_THREADS / 32; ++wv) tot += s_red[wv];
      p.loss_partial[blockIdx.x] = tot;
    }
  }
}

// Sums the per-block partials (reproducibly: the order depends only on the grid). A block owns 32 dW
// elements; its 8 warps each add a strided eighth of the partials with all loads in flight, then the
// eight subtotals are added in warp order. The last block also reduces db and the loss.
__global__ void __launch_bounds__(256)
mlp_head_reduce_kernel(const float* __restrict__ dw_partial, const float* __restrict__ db_partial,
                       const float* __restrict__ loss_partial, int nblocks, int CK, int C, int rows, float* __restrict__ dw,
                       float* __restrict__ db, float* __restrict__ loss, const float* __restrict__ dbh_partial, int K,
                       float* __restrict__ dbh) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float s_part[8][32];
  {
    // elements [0, CK) are dW, elements [CK, CK + K) are the bias gradient of the layer below (when requested)
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int i = blockIdx.x * 32 + lane;
    const int total = CK + (dbh_partial != nullptr ? K : 0);
    const bool second = i >= CK;
    const float* src = second ? dbh_partial : dw_partial;
    float* dst = second ? dbh : dw;
    const int stride = second ? K : CK;
    const bool live = i < total;
    if (second) i -= CK;
    float acc = 0.f;
    if (live) {
      for (int b0 = warp; b0 < nblocks; b0 += 8 * 8) {
        float v[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int b = b0 + 8 * u;
          v[u] = b < nblocks ? __ldcg(src + static_cast<size_t>(b) * stride + i) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) acc += v[u];
      }
    }
    s_part[warp][lane] = acc;
    __syncthreads();
    if (warp == 0 && live) {
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) tot += s_part[w][lane];
      dst[i] = tot;
    }
  }
  if (blockIdx.x == gridDim.x - 1) {
    // db and the loss: a warp per class, lanes stride over the block partials (loads in flight together),
    // then a fixed xor tree -> reproducible
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int c = warp; c <= C; c += 8) {             // c == C is the loss
      float acc = 0.f;
      for (int b = lane; b < nblocks; b += 32) acc += (c < C) ? __ldcg(db_partial + b * 16 + c) : __ldcg(loss_partial + b);
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) {
        if (c < C) db[c] = acc;
        else *loss = acc / static_cast<float>(rows);
      }
    }
  }
}

static size_t head_smem_bytes(int C, int K) {
  return (static_cast<size_t>(HD_ROWS + C) * (K + HD_PAD) + HD_ROWS * HD_DZ_LD + 16) * sizeof(float);
}

}  // namespace rtdc

using namespace rtdc;

static int head_grid(int rows) {
  const int tiles = (rows + HD_ROWS - 1) / HD_ROWS;
  int sms = 0, dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || sms <= 0) sms = 148;
  const int cap = sms * HD_BLOCKS_PER_SM;
  return tiles < cap ? tiles : cap;
}

RTDC_API size_t rtdc_mlp_head_scratch_bytes(int rows, int K, int classes) {
  const size_t g = static_cast<size_t>(head_grid(rows));
  return g * (static_cast<size_t>(classes) * K + K + 16 + 4) * sizeof(float) + 256;
}

// part: 0 = both kernels on `stream`; 1 = the pass over h only (per-block partials stay in scratch); 2 = the reduction of
// those partials only. 1 and 2 let a caller put the small reduction on a side stream, next to whatever follows the head.
static int head_train_impl(const float* h, int ldh, const float* w, const float* bias, const int64_t* labels, int rows,
                           int K, int classes, float grad_scale, float mask_scale, float* dh, int lddh, float* dw,
                           float* db, float* loss, float* db_below, void* scratch, size_t scratch_bytes, void* stream, int part) {
  if (!h || !w || !labels || !dh || !dw || !db || !loss) return fail(RTDC_ERR_ARG, "mlp_head_train: null pointer");
  if (rows <= 0 || K <= 0 || K > HD_MAXK || (K & 3) || (ldh & 3) || (lddh & 3))
    return fail(RTDC_ERR_LIMIT, "mlp_head_train: hidden width must be a multiple of 4 and <= %d, row pitches multiples of 4 (K=%d)", HD_MAXK, K);
  if (classes != 10 && classes != 16) return fail(RTDC_ERR_LIMIT, "mlp_head_train is built for 10 or 16 classes (got %d)", classes);
  if ((reinterpret_cast<uintptr_t>(h) | reinterpret_cast<uintptr_t>(w) | reinterpret_cast<uintptr_t>(dh)) & 15u)
    return fail(RTDC_ERR_ARG, "mlp_head_train: h, w and dh must be 16-byte aligned");
  const size_t need = rtdc_mlp_head_scratch_bytes(rows, K, classes);
  if (!scratch || scratch_bytes < need) return fail(RTDC_ERR_ARG, "mlp_head_train needs %zu bytes of scratch (got %zu)", need, scratch_bytes);
  const int grid = head_grid(rows);
  HeadParams p;
  p.h = h; p.ldh = ldh; p.w = w; p.bias = bias; p.labels = reinterpret_cast<const long long*>(labels);
  p.rows = rows; p.K = K; p.grad_scale = grad_scale; p.mask_scale = mask_scale; p.dh = dh; p.lddh = lddh;
  float* base = reinterpret_cast<float*>((reinterpret_cast<uintptr_t>(scratch) + 255) & ~static_cast<uintptr_t>(255));
  p.dw_partial = base;
  p.db_partial = base + static_cast<size_t>(grid) * classes * K;
  p.loss_partial = p.db_partial + static_cast<size_t>(grid) * 16;
  p.dbh_partial = db_below ? p.loss_partial + ((grid + 3) & ~3) : nullptr;
  const size_t smem = head_smem_bytes(classes, K);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (part != 2) {
    if (classes == 10) {
      RTDC_CUDA_CHECK(cudaFuncSetAttribute(mlp_head_train_kernel<10>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
      RTDC_CUDA_CHECK(launch_pdl(mlp_head_train_kernel<10>, dim3(grid), dim3(HD_THREADS), smem, st, p));
    } else {
      RTDC_CUDA_CHECK(cudaFuncSetAttribute(mlp_head_train_kernel<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem)));
      RTDC_CUDA_CHECK(launch_pdl(mlp_head_train_kernel<16>, dim3(grid), dim3(HD_THREADS), smem, st, p));
    }
    RTDC_CUDA_CHECK(cudaGetLastError());
  }
  if (part != 1) {
    const int CK = classes * K;
    const int total = CK + (db_below ? K : 0);
    RTDC_CUDA_CHECK(launch_pdl(mlp_head_reduce_kernel, dim3((total + 31) / 32), dim3(256), 0, st, p.dw_partial, p.db_partial, p.loss_partial, grid, CK,
                               classes, rows, dw, db, loss, p.dbh_partial, K, db_below));
  }
  return RTDC_OK;
}

RTDC_API int rtdc_mlp_head_train(const float* h, int ldh, const float* w, const float* bias, const int64_t* labels, int rows,
                                 int K, int classes, float grad_scale, float mask_scale, float* dh, int lddh, float* dw,
                                 float* db, float* loss, float* db_below, void* scratch, size_t scratch_bytes, void* stream) {
  return head_train_impl(h, ldh, w, bias, labels, rows, K, classes, grad_scale, mask_scale, dh, lddh, dw, db, loss, db_below, scratch,
                         scratch_bytes, stream, 0);
}

RTDC_API int rtdc_mlp_head_train_part(const float* h, int ldh, const float* w, const float* bias, const int64_t* labels, int rows,
                                      int K, int classes, float grad_scale, float mask_scale, float* dh, int lddh, float* dw,
                                      float* db, float* loss, float* db_below, void* scratch, size_t scratch_bytes, void* stream,
                                      int part) {
  if (part < 0 || part > 2) return fail(RTDC_ERR_ARG, "mlp_head_train_part: part must be 0, 1 or 2");
  return head_train_impl(h, ldh, w, bias, labels, rows, K, classes, grad_scale, mask_scale, dh, lddh, dw, db, loss, db_below, scratch,
                         scratch_bytes, stream, part);
}
