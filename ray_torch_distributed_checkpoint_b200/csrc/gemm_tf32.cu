// fp32-in / fp32-out GEMM on tcgen05 tensor cores with a fused epilogue. One CTA = one 128 x bn
// output tile (optionally one K split of it):
//
//   warp 0      TMA producer   cp.async.bulk.tensor -> 128B-swizzled smem ring, mbarrier tx-count
//   warp 1      MMA issuer     tcgen05.mma kind::tf32, fp32 accumulators in TMEM; also owns TMEM alloc
//   warps 2-5   "splitters" during the main loop, epilogue afterwards
//
// 3xTF32: the tensor core reads only the top 19 bits of each fp32 operand word. To keep fp32 GEMM
// accuracy (the reference trains in fp32 with no autocast, my_ray_module.py:141-160) the splitter
// warps compute, for every TMA-landed tile, tail = x - head(x) into a second smem tile of the SAME
// swizzled layout (an elementwise pass, so no layout maths), and the issuer runs three MMAs per
// K step: tail(A)*B + A*tail(B) + A*B. HBM/L2 traffic stays that of a plain fp32 GEMM.
//
// Operand layouts (both supported for A and for B, chosen by the tensor maps + descriptor bits):
//   K-major : row-major [rows, K]   -> one TMA box of [rows][32 floats]
//   MN-major: row-major [K, rows]   -> rows/32 TMA boxes of [32 k][32 floats]
// so forward (X.W^T), dgrad (dY.W) and wgrad (dY^T.X) all run on this kernel with no transposes
// in memory (K1-K3, K6, K12 in SURVEY.md §2.5).
#include "common.h"
#include "ptx.cuh"
#include "../../include/rtdc_b200.h"

namespace rtdc {

constexpr int BM = 128;           // tile rows == TMEM lanes
constexpr int BK = 32;            // fp32 elements per K block == one 128-byte swizzle row
constexpr int UMMA_K = 8;         // tf32: 32 bytes per instruction along K
constexpr int MAX_STAGES = 4;
constexpr int MAX_BN = 256;
constexpr int TMEM_COLS = 256;
constexpr int NUM_THREADS = 192;  // 6 warps
constexpr int EPI_THREADS = 128;  // warps 2..5
constexpr int SMEM_LIMIT = 232448;  // 227 KB opt-in maximum per CTA

struct GemmParams {
  int M, N, K;
  int bn, stages, splits, kb_per_split, num_kb;
  int m_tiles, n_tiles;
  uint32_t a_bytes, b_bytes, stage_bytes;  // per stage: raw A, raw B, total (incl. tails when split)
  float* out; int ldo; int n_store;
  const float* bias; int relu;
  float drop_scale; uint32_t drop_thr16; uint32_t seed_lo, seed_hi; const uint32_t* step_ptr; uint32_t step_imm, salt;
  const float* mask_src; int ldm; float mask_scale;
  float* scratch; unsigned int* counters;
};

// keep flag for dropout: 16 random bits per element, 8 elements per Philox call.
__device__ __forceinline__ uint32_t rand16(const Philox4& r, int j) {
  const uint32_t w = (j < 2) ? r.x : (j < 4) ? r.y : (j < 6) ? r.z : r.w;
  return (j & 1) ? (w >> 16) : (w & 0xFFFFu);
}

template <bool A_MN, bool B_MN, int NSPLIT>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmParams p) {
  extern __shared__ uint8_t smem_dyn[];
  __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t ready_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t accum_bar;
  __shared__ uint32_t tmem_base_slot;
  __shared__ uint32_t last_flag;

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;

  // 1024-byte aligned stage ring (128B swizzle atoms are 1024 B)
  const uint32_t dyn_base = smem_u32(smem_dyn);
  const uint32_t ring = (dyn_base + 1023u) & ~1023u;
  uint8_t* ring_ptr = smem_dyn + (ring - dyn_base);

  const int tile = blockIdx.x;
  const int m_tile = tile % p.m_tiles;
  const int n_tile = tile / p.m_tiles;
  const int m0 = m_tile * BM;
  const int n0 = n_tile * p.bn;
  const int split = blockIdx.y;
  const int kb_begin = split * p.kb_per_split;
  const int kb_end = min(p.num_kb, kb_begin + p.kb_per_split);
  const int n_iters = kb_end - kb_begin;

  // columns this tile really owns, and the (multiple of 16) N the MMA runs at
  const int n_valid = min(p.bn, p.N - n0);
  const uint32_t umma_n = static_cast<uint32_t>(min(p.bn, (n_valid + 15) & ~15));

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&ready_bar[s], EPI_THREADS / 32);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(&accum_bar, 1);
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
  }
  if (warp == 1) {
    tmem_alloc(&tmem_base_slot, TMEM_COLS);
    tmem_relinquish();
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = tmem_base_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      for (int it = 0; it < n_iters; ++it) {
        const int s = it % p.stages;
        const uint32_t ph = (it / p.stages) & 1;
        mbar_wait(&empty_bar[s], ph ^ 1u);
        uint8_t* sa = ring_ptr + static_cast<size_t>(s) * p.stage_bytes;
        uint8_t* sb = sa + p.a_bytes;
        mbar_expect_tx(&full_bar[s], p.a_bytes + p.b_bytes);
        const int k0 = (kb_begin + it) * BK;
        if (!A_MN) {
          tma_load_2d(sa, &tmA, &full_bar[s], k0, m0);
        } else {
#pragma unroll
          for (int b = 0; b < BM / 32; ++b) tma_load_2d(sa + b * 4096, &tmA, &full_bar[s], m0 + 32 * b, k0);
        }
        if (!B_MN) {
          tma_load_2d(sb, &tmB, &full_bar[s], k0, n0);
        } else {
          for (int b = 0; b < p.bn / 32; ++b) tma_load_2d(sb + b * 4096, &tmB, &full_bar[s], n0 + 32 * b, k0);
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      const uint32_t idesc = umma_idesc(UMMA_TF32, A_MN, B_MN, umma_n);
      const uint32_t tail_off = p.a_bytes + p.b_bytes;  // tails live right after the raw tiles
      // K-major: 8-row groups 1024 B apart (SW128). MN-major: 32-wide MN blocks 4096 B apart, 4-row
      // K groups 512 B apart (SW128 with 32-byte atoms: the only MN-major layout tf32 accepts).
      const uint32_t a_lbo = A_MN ? 4096u : 16u, b_lbo = B_MN ? 4096u : 16u;
      const uint32_t a_sbo = A_MN ? 512u : 1024u, b_sbo = B_MN ? 512u : 1024u;
      const uint32_t a_lt = A_MN ? UMMA_LAYOUT_SW128_BASE32B : UMMA_LAYOUT_SW128;
      const uint32_t b_lt = B_MN ? UMMA_LAYOUT_SW128_BASE32B : UMMA_LAYOUT_SW128;
      const uint32_t a_kstep = A_MN ? 1024u : 32u, b_kstep = B_MN ? 1024u : 32u;
      for (int it = 0; it < n_iters; ++it) {
        const int s = it % p.stages;
        const uint32_t ph = (it / p.stages) & 1;
        mbar_wait(NSPLIT == 3 ? &ready_bar[s] : &full_bar[s], ph);
        tcgen05_fence_after();
        const uint32_t sa = ring + static_cast<uint32_t>(s) * p.stage_bytes;
        const uint32_t sb = sa + p.a_bytes;
        const int k_left = p.K - (kb_begin + it) * BK;
        const int ksteps = min(BK / UMMA_K, (k_left + UMMA_K - 1) / UMMA_K);
        for (int j = 0; j < ksteps; ++j) {
          const uint64_t da = umma_smem_desc(sa + j * a_kstep, a_lbo, a_sbo, a_lt);
          const uint64_t db = umma_smem_desc(sb + j * b_kstep, b_lbo, b_sbo, b_lt);
          const uint32_t acc = (it > 0 || j > 0) ? 1u : 0u;
          if (NSPLIT == 3) {
            const uint64_t da_t = umma_smem_desc(sa + tail_off + j * a_kstep, a_lbo, a_sbo, a_lt);
            const uint64_t db_t = umma_smem_desc(sb + tail_off + j * b_kstep, b_lbo, b_sbo, b_lt);
            umma_tf32_ss(tmem_base, da_t, db, idesc, acc);  // small terms first
            umma_tf32_ss(tmem_base, da, db_t, idesc, 1u);
            umma_tf32_ss(tmem_base, da, db, idesc, 1u);
          } else {
            umma_tf32_ss(tmem_base, da, db, idesc, acc);
          }
        }
        tcgen05_commit(&empty_bar[s]);  // smem slot reusable once these MMAs have read it
      }
      tcgen05_commit(&accum_bar);       // accumulators complete
    }
    __syncwarp();
  } else {
    // ------------------------------------------------------------------ splitter, then epilogue
    const int et = threadIdx.x - 64;  // 0..127
    if (NSPLIT == 3) {
      const uint32_t raw_bytes = p.a_bytes + p.b_bytes;
      const int nvec = static_cast<int>(raw_bytes >> 4);
      for (int it = 0; it < n_iters; ++it) {
        const int s = it % p.stages;
        const uint32_t ph = (it / p.stages) & 1;
        mbar_wait(&full_bar[s], ph);
        const float4* raw = reinterpret_cast<const float4*>(ring_ptr + static_cast<size_t>(s) * p.stage_bytes);
        float4* tail = reinterpret_cast<float4*>(ring_ptr + static_cast<size_t>(s) * p.stage_bytes + raw_bytes);
#pragma unroll 4
        for (int i = et; i < nvec; i += EPI_THREADS) {
          const float4 v = raw[i];
          float4 t;
          t.x = v.x - __uint_as_float(__float_as_uint(v.x) & 0xFFFFE000u);
          t.y = v.y - __uint_as_float(__float_as_uint(v.y) & 0xFFFFE000u);
          t.z = v.z - __uint_as_float(__float_as_uint(v.z) & 0xFFFFE000u);
          t.w = v.w - __uint_as_float(__float_as_uint(v.w) & 0xFFFFE000u);
          tail[i] = t;
        }
        fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's async proxy
        __syncwarp();
        if (lane == 0) mbar_arrive(&ready_bar[s]);
      }
    }

    mbar_wait(&accum_bar, 0);
    tcgen05_fence_after();

    const int wq = warp & 3;                 // TMEM lane quarter this warp may read
    const int r_in_tile = wq * 32 + lane;
    const int row = m0 + r_in_tile;
    const uint32_t taddr = tmem_base + (static_cast<uint32_t>(wq * 32) << 16);
    const int n_cols_out = min(p.bn, p.n_store - n0);  // columns to write (valid + zero padding)
    const int n_chunks = (n_cols_out + 31) / 32;
    const int acc_chunks = (static_cast<int>(umma_n) + 31) / 32;  // chunks that hold MMA output
    uint32_t r[32];

    bool finalize = true;
    if (p.splits > 1) {
      // write this split's partial tile; the last CTA to arrive for the tile sums them in split order
      float* my = p.scratch + (static_cast<size_t>(tile) * p.splits + split) * (BM * MAX_BN);
      for (int c = 0; c < acc_chunks; ++c) {
        tmem_ld_32x32(taddr + c * 32, r);
        tmem_ld_wait();
        float4* dst = reinterpret_cast<float4*>(my + (static_cast<size_t>(c) * BM + r_in_tile) * 32);
#pragma unroll
        for (int q = 0; q < 8; ++q)
          dst[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                               __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
      }
      __threadfence();
      named_bar_sync(1, EPI_THREADS);
      if (et == 0) {
        const unsigned int old = atomicAdd(&p.counters[tile], 1u);
        last_flag = (old == static_cast<unsigned int>(p.splits - 1)) ? 1u : 0u;
        if (last_flag) p.counters[tile] = 0u;  // leave the counter ready for the next launch
      }
      named_bar_sync(1, EPI_THREADS);
      finalize = last_flag != 0u;
      if (finalize) __threadfence();
    }

    if (finalize) {
      uint32_t step = p.step_imm;
      if (p.drop_thr16 != 0u && p.step_ptr != nullptr) step = *p.step_ptr;
      for (int c = 0; c < n_chunks; ++c) {
        float v[32];
        if (p.splits > 1) {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
          if (c < acc_chunks) {
            for (int z = 0; z < p.splits; ++z) {
              const float4* src = reinterpret_cast<const float4*>(
                  p.scratch + (static_cast<size_t>(tile) * p.splits + z) * (BM * MAX_BN) +
                  (static_cast<size_t>(c) * BM + r_in_tile) * 32);
#pragma unroll
              for (int q = 0; q < 8; ++q) {
                const float4 t = __ldcg(src + q);
                v[4 * q] += t.x; v[4 * q + 1] += t.y; v[4 * q + 2] += t.z; v[4 * q + 3] += t.w;
              }
            }
          }
        } else {
          if (c < acc_chunks) {
            tmem_ld_32x32(taddr + c * 32, r);
            tmem_ld_wait();
          }
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = (c < acc_chunks) ? __uint_as_float(r[j]) : 0.f;
        }

        const int col0 = n0 + c * 32;
        // bias + ReLU
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const int col = col0 + j;
          float x = v[j];
          if (col < p.N) {
            if (p.bias != nullptr) x += __ldg(p.bias + col);
            if (p.relu) x = fmaxf(x, 0.f);
          } else {
            x = 0.f;
          }
          v[j] = x;
        }
        // dropout: element index e = row*N + col, 8 elements per Philox call, 16 bits each
        if (p.drop_thr16 != 0u && row < p.M) {
#pragma unroll
          for (int g = 0; g < 4; ++g) {
            const unsigned long long e0 =
                static_cast<unsigned long long>(row) * static_cast<unsigned long long>(p.N) + (col0 + g * 8);
            if ((e0 & 7ull) == 0ull) {
              const unsigned long long grp = e0 >> 3;
              const Philox4 rnd = philox4x32_10(static_cast<uint32_t>(grp), static_cast<uint32_t>(grp >> 32), step,
                                                p.salt, p.seed_lo, p.seed_hi);
#pragma unroll
              for (int j = 0; j < 8; ++j)
                v[g * 8 + j] = (rand16(rnd, j) >= p.drop_thr16) ? v[g * 8 + j] * p.drop_scale : 0.f;
            } else {
              // unaligned row start (N not a multiple of 8): one Philox call per element group it touches
              for (int j = 0; j < 8; ++j) {
                const unsigned long long e = e0 + j;
                const unsigned long long grp = e >> 3;
                const Philox4 rnd = philox4x32_10(static_cast<uint32_t>(grp), static_cast<uint32_t>(grp >> 32),
                                                  step, p.salt, p.seed_lo, p.seed_hi);
                v[g * 8 + j] = (rand16(rnd, static_cast<int>(e & 7ull)) >= p.drop_thr16) ? v[g * 8 + j] * p.drop_scale
                                                                                       : 0.f;
              }
            }
          }
        }
        // backward mask of the previous layer's ReLU+dropout output
        if (p.mask_src != nullptr && row < p.M) {
          const float* ms = p.mask_src + static_cast<size_t>(row) * p.ldm + col0;
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            if (col0 + j < p.N) v[j] = (__ldg(ms + j) > 0.f) ? v[j] * p.mask_scale : 0.f;
          }
        }
        if (row < p.M) {
          float* dst = p.out + static_cast<size_t>(row) * p.ldo + col0;
          const int rem = n_cols_out - c * 32;
          if (rem >= 32 && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0)) {
#pragma unroll
            for (int q = 0; q < 8; ++q)
              reinterpret_cast<float4*>(dst)[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j)
              if (j < rem) dst[j] = v[j];
          }
        }
      }
    }
  }

  // teardown: all TMEM reads done before the allocating warp frees the columns
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------ host side

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  if (fn) return fn;
  void* sym = nullptr;
  cudaDriverEntryPointQueryResult q;
  if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess ||
      q != cudaDriverEntryPointSuccess)
    return nullptr;
  fn = reinterpret_cast<EncodeTiledFn>(sym);
  return fn;
}

// 2-D fp32 tensor map over a row-major [outer, inner] matrix with leading dimension ld (elements).
static int make_map(CUtensorMap* map, const float* base, int inner, int outer, int ld, int box_inner, int box_outer,
                    bool mn_major) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return fail(RTDC_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  if ((reinterpret_cast<uintptr_t>(base) & 15u) != 0) return fail(RTDC_ERR_ARG, "operand pointer must be 16-byte aligned");
  if ((static_cast<size_t>(ld) * 4) % 16 != 0) return fail(RTDC_ERR_ARG, "operand leading dimension must be a multiple of 4 floats (got %d)", ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 4};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, mn_major ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(RTDC_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  return RTDC_OK;
}

struct Plan {
  int bn, stages, splits, kb_per_split, num_kb, m_tiles, n_tiles;
  uint32_t a_bytes, b_bytes, stage_bytes;
  size_t smem;
};

static int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n <= 0)
      n = 148;
  }
  return n;
}

static int make_plan(const rtdc_gemm_desc* d, Plan* pl, bool need_device) {
  if (!d) return fail(RTDC_ERR_ARG, "null descriptor");
  if (d->M <= 0 || d->N <= 0 || d->K <= 0) return fail(RTDC_ERR_ARG, "M, N, K must be positive (got %d, %d, %d)", d->M, d->N, d->K);
  if (d->precision != 1 && d->precision != 3) return fail(RTDC_ERR_ARG, "precision must be 1 (tf32) or 3 (3xTF32)");
  const int gran = d->b_mn ? 32 : 16;
  int bn = d->bn;
  if (bn == 0) {
    const int nt = ceil_div(d->N, MAX_BN);
    bn = round_up(ceil_div(d->N, nt), gran);
  }
  if (bn < gran || bn > MAX_BN || bn % gran != 0)
    return fail(RTDC_ERR_ARG, "bn must be a multiple of %d in [%d, %d] (got %d)", gran, gran, MAX_BN, bn);
  pl->bn = bn;
  pl->m_tiles = ceil_div(d->M, BM);
  pl->n_tiles = ceil_div(d->N, bn);
  if (d->n_store > d->N && d->n_store > pl->n_tiles * bn)
    return fail(RTDC_ERR_ARG, "n_store (%d) reaches beyond the last N tile", d->n_store);
  pl->num_kb = ceil_div(d->K, BK);
  pl->a_bytes = BM * BK * 4;
  pl->b_bytes = static_cast<uint32_t>(bn) * BK * 4;
  pl->stage_bytes = (pl->a_bytes + pl->b_bytes) * (d->precision == 3 ? 2u : 1u);
  int stages = static_cast<int>((SMEM_LIMIT - 1024 - 256) / pl->stage_bytes);
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages > pl->num_kb) stages = pl->num_kb < 1 ? 1 : pl->num_kb;
  if (stages < 1) return fail(RTDC_ERR_LIMIT, "tile does not fit shared memory");
  pl->stages = stages;
  pl->smem = static_cast<size_t>(stages) * pl->stage_bytes + 1024;
  const int tiles = pl->m_tiles * pl->n_tiles;
  int splits = d->splits;
  if (splits == 0) {
    splits = 1;
    const int sms = need_device ? num_sms() : 148;
    if (tiles * 2 <= sms && pl->num_kb >= 16) {
      splits = sms / tiles;
      const int max_by_k = pl->num_kb / 8;  // keep >= 8 K blocks per split
      if (splits > max_by_k) splits = max_by_k;
      if (splits > 64) splits = 64;
      if (splits < 1) splits = 1;
    }
  }
  if (splits > pl->num_kb) splits = pl->num_kb;
  pl->kb_per_split = ceil_div(pl->num_kb, splits);
  pl->splits = ceil_div(pl->num_kb, pl->kb_per_split);  // no empty split
  return RTDC_OK;
}

// scratch layout: [COUNTER_BYTES of per-tile arrival counters][tiles x splits partial tiles]. The
// counters sit at a fixed place so one buffer can serve differently shaped calls one after another.
constexpr size_t COUNTER_BYTES = 4096;
static size_t scratch_bytes_for(const Plan& pl) {
  if (pl.splits <= 1) return 0;
  const size_t tiles = static_cast<size_t>(pl.m_tiles) * pl.n_tiles;
  return COUNTER_BYTES + tiles * pl.splits * (static_cast<size_t>(BM) * MAX_BN * 4);
}

template <bool A_MN, bool B_MN, int NSPLIT>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const GemmParams& gp, const Plan& pl, cudaStream_t st) {
  auto kern = gemm_tf32_kernel<A_MN, B_MN, NSPLIT>;
  RTDC_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT - 1024));
  dim3 grid(pl.m_tiles * pl.n_tiles, pl.splits, 1);
  kern<<<grid, NUM_THREADS, pl.smem, st>>>(ta, tb, gp);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

}  // namespace rtdc

using namespace rtdc;

RTDC_API size_t rtdc_gemm_scratch_bytes(const rtdc_gemm_desc* d) {
  Plan pl;
  if (make_plan(d, &pl, true) != RTDC_OK) return 0;
  return scratch_bytes_for(pl);
}

RTDC_API int rtdc_gemm_plan(const rtdc_gemm_desc* d, int* out5) {
  Plan pl;
  int rc = make_plan(d, &pl, true);
  if (rc != RTDC_OK) return rc;
  out5[0] = pl.bn; out5[1] = pl.stages; out5[2] = pl.splits; out5[3] = pl.m_tiles * pl.n_tiles;
  out5[4] = static_cast<int>(pl.smem);
  return RTDC_OK;
}

RTDC_API int rtdc_gemm(const rtdc_gemm_desc* d, void* stream) {
  Plan pl;
  int rc = make_plan(d, &pl, true);
  if (rc != RTDC_OK) return rc;
  if (!d->a || !d->b || !d->out) return fail(RTDC_ERR_ARG, "a, b and out must be non-null");
  if (d->drop_p < 0.f || d->drop_p >= 1.f) return fail(RTDC_ERR_ARG, "drop_p must be in [0, 1)");
  const int n_store = d->n_store > d->N ? d->n_store : d->N;
  if (d->ldo < n_store) return fail(RTDC_ERR_ARG, "ldo (%d) smaller than the stored width (%d)", d->ldo, n_store);

  CUtensorMap ta, tb;
  if (!d->a_mn) rc = make_map(&ta, d->a, d->K, d->M, d->lda, BK, BM, false);
  else          rc = make_map(&ta, d->a, d->M, d->K, d->lda, 32, BK, true);
  if (rc != RTDC_OK) return rc;
  if (!d->b_mn) rc = make_map(&tb, d->b, d->K, d->N, d->ldb, BK, pl.bn, false);
  else          rc = make_map(&tb, d->b, d->N, d->K, d->ldb, 32, BK, true);
  if (rc != RTDC_OK) return rc;

  GemmParams gp;
  memset(&gp, 0, sizeof(gp));
  gp.M = d->M; gp.N = d->N; gp.K = d->K;
  gp.bn = pl.bn; gp.stages = pl.stages; gp.splits = pl.splits; gp.kb_per_split = pl.kb_per_split; gp.num_kb = pl.num_kb;
  gp.m_tiles = pl.m_tiles; gp.n_tiles = pl.n_tiles;
  gp.a_bytes = pl.a_bytes; gp.b_bytes = pl.b_bytes; gp.stage_bytes = pl.stage_bytes;
  gp.out = d->out; gp.ldo = d->ldo; gp.n_store = n_store;
  gp.bias = d->bias; gp.relu = d->relu;
  if (d->drop_p > 0.f) {
    uint32_t thr = static_cast<uint32_t>(d->drop_p * 65536.0f + 0.5f);
    if (thr == 0) thr = 1;
    if (thr > 65535u) thr = 65535u;
    gp.drop_thr16 = thr;
    gp.drop_scale = 1.0f / (1.0f - d->drop_p);
  }
  gp.seed_lo = static_cast<uint32_t>(d->seed); gp.seed_hi = static_cast<uint32_t>(d->seed >> 32);
  gp.step_ptr = d->step_ptr; gp.step_imm = d->step_imm; gp.salt = d->salt;
  gp.mask_src = d->mask_src; gp.ldm = d->ldm; gp.mask_scale = d->mask_scale;
  if (pl.splits > 1) {
    const size_t need = scratch_bytes_for(pl);
    if (!d->scratch || d->scratch_bytes < need)
      return fail(RTDC_ERR_ARG, "split-K x%d needs %zu bytes of scratch (got %zu)", pl.splits, need, d->scratch_bytes);
    if (static_cast<size_t>(pl.m_tiles) * pl.n_tiles * 4 > COUNTER_BYTES)
      return fail(RTDC_ERR_LIMIT, "split-K supports at most %zu output tiles", COUNTER_BYTES / 4);
    gp.counters = reinterpret_cast<unsigned int*>(d->scratch);
    gp.scratch = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(d->scratch) + COUNTER_BYTES);
  }
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool three = d->precision == 3;
  if (!d->a_mn && !d->b_mn) return three ? launch<false, false, 3>(ta, tb, gp, pl, st) : launch<false, false, 1>(ta, tb, gp, pl, st);
  if (!d->a_mn && d->b_mn)  return three ? launch<false, true, 3>(ta, tb, gp, pl, st)  : launch<false, true, 1>(ta, tb, gp, pl, st);
  if (d->a_mn && !d->b_mn)  return three ? launch<true, false, 3>(ta, tb, gp, pl, st)  : launch<true, false, 1>(ta, tb, gp, pl, st);
  return three ? launch<true, true, 3>(ta, tb, gp, pl, st) : launch<true, true, 1>(ta, tb, gp, pl, st);
}
