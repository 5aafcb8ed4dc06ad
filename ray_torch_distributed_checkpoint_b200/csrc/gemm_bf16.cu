// bf16 GEMM on the 5th-generation tensor cores for the north-star model configurations (GPT-2 small / ResNet-50 heads in
// bf16): D[M,N] = epilogue( sum_k A(m,k) B(n,k) ), bf16 operands, fp32 accumulation in TMEM, bf16 or fp32 output.
//
//   tcgen05.mma.cta_group::1.kind::f16 (M = 128, N <= 256, K = 16 per instruction), one elected thread issues;
//   operands staged by TMA (cp.async.bulk.tensor, 128-byte swizzle) through a 4-stage mbarrier ring: 64 K-elements
//   (= 128 bytes) per stage; accumulators in two 256-column TMEM buffers, so the epilogue of tile j (tcgen05.ld ->
//   alpha, bias, activation -> global) overlaps the main loop of tile j+1; persistent CTAs walk (tile x K-split) units.
//
// Both operand majors on one kernel, so a Linear layer's forward (x W^T), dgrad (dy W) and wgrad (dy^T x) read the
// tensors as they are stored (no transposed copies):
//   K-major  operand  X[rows, K]  row-major: TMA box {64 k, rows}; smem rows of 128 B, 8-row groups 1024 B apart.
//   MN-major operand  X[K, rows]  row-major: TMA boxes {64 rows, 64 k}; smem = per 64-row block 64 K-rows of 128 B
//                     (canonical SW128 MN-major layout for 16-bit types: 8-K-row groups 1024 B apart = SBO, next 64-row
//                     block one box = 8192 B further = LBO, one K = 16 instruction consumes 2048 B of a block).
// Split-K (wgrad: K = tokens of the batch) accumulates with fp32 atomics into a zeroed fp32 output.
//
// The reference trains its MLP in fp32 (the tf32x3 kernels of gemm_tf32.cu cover it); this file has no counterpart there.
#include <cuda_bf16.h>

#include "common.h"
#include "ptx.cuh"
#include "../../include/rtdc_b200.h"

namespace rtdc {
namespace bf16gemm {

constexpr int BM = 128;
constexpr int BK = 64;              // bf16 elements per stage along K = 128 bytes = one swizzle span
constexpr int UMMA_K = 16;          // kind::f16: 32 bytes of K per instruction
constexpr int MAX_BN = 256;
constexpr int STAGES = 4;
constexpr int TMEM_COLS = 512;      // two accumulator buffers of MAX_BN fp32 columns
constexpr int NUM_THREADS = 192;    // warp 0 TMA producer, warp 1 MMA issuer + TMEM owner, warps 2-5 epilogue
constexpr int A_BYTES = BM * BK * 2;             // 16 KB
constexpr int BOX_BYTES = 64 * BK * 2;           // one MN-major box: 64 rows x 64 k = 8 KB

struct Params {
  int M, N, K;
  int bn, m_tiles, n_tiles, splits, num_kb, kb_per_split, total_units;
  uint32_t b_bytes, stage_bytes;
  void* out; int ldo; int out_fp32;
  const void* bias; int bias_fp32;
  int act;            // 0 none, 1 relu, 2 gelu (tanh approximation), 3 gelu (erf)
  float alpha;
  int tma_store;      // CTA-pair kernel: bf16 output leaves through swizzled smem boxes + TMA stores
};

__device__ __forceinline__ float apply_act(float x, int act) {
  if (act == 1) return fmaxf(x, 0.f);
  if (act == 2) {
    const float u = 0.7978845608028654f * (x + 0.044715f * x * x * x);
    return 0.5f * x * (1.f + tanhf(u));
  }
  if (act == 3) return 0.5f * x * (1.f + erff(x * 0.7071067811865476f));
  return x;
}

// One epilogue warp's share of a finished accumulator tile: its 32 TMEM lanes (rows), all valid columns, 32 at a time with the
// next chunk's tcgen05.ld in flight while the current one is scaled, biased, activated and stored.
__device__ __forceinline__ void epilogue_tile(const Params& p, uint32_t taddr, int row, int n0, int n_valid, bool has_bias,
                                              const float* bias_s, int c_begin = 0, int c_stride = 1) {
  // chunks c_begin, c_begin + c_stride, ... (two warps share a TMEM lane quarter in the CTA-pair kernel and interleave)
  const int n_chunks = (n_valid + 31) / 32;
  uint32_t r[32];
  if (c_begin < n_chunks) tmem_ld_32x32(taddr + c_begin * 32, r);
  for (int c = c_begin; c < n_chunks; c += c_stride) {
    tmem_ld_wait();
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha;
    if (c + c_stride < n_chunks) tmem_ld_32x32(taddr + (c + c_stride) * 32, r);     // in flight while this chunk is processed
    const int col0 = n0 + c * 32;
    if (has_bias) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float4 bq = *reinterpret_cast<const float4*>(&bias_s[c * 32 + 4 * i]);
        v[4 * i] += bq.x; v[4 * i + 1] += bq.y; v[4 * i + 2] += bq.z; v[4 * i + 3] += bq.w;
      }
    }
    if (p.act != 0) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = apply_act(v[i], p.act);
    }
    if (row < p.M) {
      const int ncol = min(32, p.N - col0);
      if (p.out_fp32) {
        float* dst = static_cast<float*>(p.out) + static_cast<size_t>(row) * p.ldo + col0;
        const bool vec = ncol == 32 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0;
        if (p.splits > 1) {
          if (vec) {
#pragma unroll
            for (int i = 0; i < 8; ++i)
              asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(dst + 4 * i), "f"(v[4 * i]), "f"(v[4 * i + 1]),
                           "f"(v[4 * i + 2]), "f"(v[4 * i + 3]) : "memory");
          } else {
#pragma unroll
            for (int i = 0; i < 32; ++i)
              if (i < ncol) atomicAdd(dst + i, v[i]);
          }
        } else if (vec) {
#pragma unroll
          for (int i = 0; i < 8; ++i) reinterpret_cast<float4*>(dst)[i] = make_float4(v[4 * i], v[4 * i + 1], v[4 * i + 2], v[4 * i + 3]);
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (i < ncol) dst[i] = v[i];
        }
      } else {
        __nv_bfloat16* dst = static_cast<__nv_bfloat16*>(p.out) + static_cast<size_t>(row) * p.ldo + col0;
        if (ncol == 32 && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0) {
#pragma unroll
          for (int i = 0; i < 4; ++i) {
            uint32_t w[4];
#pragma unroll
            for (int t = 0; t < 4; ++t) {
              const __nv_bfloat162 h = __floats2bfloat162_rn(v[8 * i + 2 * t], v[8 * i + 2 * t + 1]);
              w[t] = *reinterpret_cast<const uint32_t*>(&h);
            }
            reinterpret_cast<uint4*>(dst)[i] = make_uint4(w[0], w[1], w[2], w[3]);
          }
        } else {
#pragma unroll
          for (int i = 0; i < 32; ++i)
            if (i < ncol) dst[i] = __float2bfloat16_rn(v[i]);
        }
      }
    }
  }
}

// bf16 output through shared memory and TMA: a warp packs its 32 x 32 chunk into a 64-byte-swizzled 2 KB box (conflict-free:
// each lane owns a row) and one lane stores the box with cp.async.bulk.tensor — full 64-byte row segments per request instead
// of 32 scattered 16-byte writes per store instruction, which is what bound the direct-store epilogue on the forward shapes.
// Two boxes per warp, so packing chunk c+1 overlaps the store of chunk c. TMA clips rows >= M and columns >= N.
__device__ __forceinline__ void epilogue_tile_tma(const Params& p, const CUtensorMap* tmOut, uint32_t taddr, int row0, int n0, int n_valid,
                                                  bool has_bias, const float* bias_s, unsigned char* stage2, int lane, int c_begin,
                                                  int c_stride) {
  const int n_chunks = (n_valid + 31) / 32;
  uint32_t r[32];
  if (c_begin < n_chunks) tmem_ld_32x32(taddr + c_begin * 32, r);
  int parity = 0;
  for (int c = c_begin; c < n_chunks; c += c_stride, parity ^= 1) {
    tmem_ld_wait();
    float v[32];
#pragma unroll
    for (int i = 0; i < 32; ++i) v[i] = __uint_as_float(r[i]) * p.alpha;
    if (c + c_stride < n_chunks) tmem_ld_32x32(taddr + (c + c_stride) * 32, r);
    if (has_bias) {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float4 bq = *reinterpret_cast<const float4*>(&bias_s[c * 32 + 4 * i]);
        v[4 * i] += bq.x; v[4 * i + 1] += bq.y; v[4 * i + 2] += bq.z; v[4 * i + 3] += bq.w;
      }
    }
    if (p.act != 0) {
#pragma unroll
      for (int i = 0; i < 32; ++i) v[i] = apply_act(v[i], p.act);
    }
    unsigned char* box = stage2 + parity * 2048;
    if (lane == 0) tma_store_wait_read<1>();          // the store that last read this box (two chunks ago) has drained it
    __syncwarp();
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      uint32_t w[4];
#pragma unroll
      for (int t = 0; t < 4; ++t) {
        const __nv_bfloat162 h = __floats2bfloat162_rn(v[8 * i + 2 * t], v[8 * i + 2 * t + 1]);
        w[t] = *reinterpret_cast<const uint32_t*>(&h);
      }
      // CU_TENSOR_MAP_SWIZZLE_64B: 16-byte chunk index XOR address bits 7-8, i.e. (row >> 1) & 3 with 64-byte rows
      *reinterpret_cast<uint4*>(box + lane * 64 + ((i ^ ((lane >> 1) & 3)) << 4)) = make_uint4(w[0], w[1], w[2], w[3]);
    }
    fence_proxy_async_smem();
    __syncwarp();
    if (lane == 0) {
      tma_store_2d(tmOut, box, n0 + c * 32, row0);
      tma_store_commit();
    }
  }
}

template <bool A_MN, bool B_MN>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const Params p) {
  pdl_launch_dependents();
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[STAGES], empty_bar[STAGES], tmem_full_bar[2], tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) float bias_s[MAX_BN];

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);   // warp-uniform for the compiler
  const int lane = threadIdx.x & 31;
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full_bar[b], 1); mbar_init(&tmem_empty_bar[b], 4); }
    mbar_fence_init();
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
  pdl_wait();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_slot, 0);
  const int total_units = p.total_units;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (converged warp, one lane issues)
    int s = 0;
    uint32_t ph = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
      const int split = u % p.splits, tile = u / p.splits;
      const int m0 = (tile % p.m_tiles) * BM, n0 = (tile / p.m_tiles) * p.bn;
      const int kb0 = split * p.kb_per_split, kb1 = min(p.num_kb, kb0 + p.kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[s], ph ^ 1u);
        __syncwarp();
        if (elect_one()) {
          const uint32_t sa = ring + static_cast<uint32_t>(s) * p.stage_bytes, sb = sa + A_BYTES;
          mbar_expect_tx(&full_bar[s], p.stage_bytes);
          const int k0 = kb * BK;
          if (A_MN) {
#pragma unroll
            for (int b = 0; b < BM / 64; ++b)
              tma_load_2d(reinterpret_cast<void*>(__cvta_shared_to_generic(sa + b * BOX_BYTES)), &tmA, &full_bar[s], m0 + b * 64, k0);
          } else {
            tma_load_2d(reinterpret_cast<void*>(__cvta_shared_to_generic(sa)), &tmA, &full_bar[s], k0, m0);
          }
          if (B_MN) {
            for (int b = 0; b < p.bn / 64; ++b)
              tma_load_2d(reinterpret_cast<void*>(__cvta_shared_to_generic(sb + b * BOX_BYTES)), &tmB, &full_bar[s], n0 + b * 64, k0);
          } else {
            tma_load_2d(reinterpret_cast<void*>(__cvta_shared_to_generic(sb)), &tmB, &full_bar[s], k0, n0);
          }
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    constexpr uint32_t a_lbo = A_MN ? BOX_BYTES : 16u, b_lbo = B_MN ? BOX_BYTES : 16u;
    constexpr uint32_t sbo = 1024u;                       // 8 rows (K-major) / 8 K-rows (MN-major) of 128 B
    constexpr uint32_t a_kstep = A_MN ? UMMA_K * 128u : UMMA_K * 2u, b_kstep = B_MN ? UMMA_K * 128u : UMMA_K * 2u;
    const uint32_t idesc = umma_idesc(UMMA_BF16, A_MN, B_MN, static_cast<uint32_t>(p.bn));
    int s = 0;
    uint32_t ph = 0;
    int j = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x, ++j) {
      const int split = u % p.splits;
      const int kb0 = split * p.kb_per_split, kb1 = min(p.num_kb, kb0 + p.kb_per_split);
      const int buf = j & 1;
      mbar_wait(&tmem_empty_bar[buf], ((j >> 1) & 1) ^ 1u);      // the epilogue has drained this accumulator buffer
      tcgen05_fence_after();
      const uint32_t tmem_acc = tmem_base + static_cast<uint32_t>(buf) * MAX_BN;
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[s], ph);
        tcgen05_fence_after();
        __syncwarp();
        const uint32_t sa = ring + static_cast<uint32_t>(s) * p.stage_bytes, sb = sa + A_BYTES;
        if (elect_one()) {
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; ++k) {
            const uint64_t da = umma_smem_desc(sa + k * a_kstep, a_lbo, sbo, UMMA_LAYOUT_SW128);
            const uint64_t db = umma_smem_desc(sb + k * b_kstep, b_lbo, sbo, UMMA_LAYOUT_SW128);
            umma_f16_ss(tmem_acc, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
          }
          tcgen05_commit(&empty_bar[s]);                          // stage reusable once these MMAs have read it
          if (kb == kb1 - 1) tcgen05_commit(&tmem_full_bar[buf]);  // accumulator complete
        }
        __syncwarp();
        if (++s == STAGES) { s = 0; ph ^= 1u; }
      }
      if (kb1 <= kb0 && elect_one()) tcgen05_commit(&tmem_full_bar[buf]);   // empty split (cannot happen with the host's plan)
    }
  } else {
    // ------------------------------------------------------------------ epilogue: TMEM -> registers -> global
    // The four warps first stage the tile's bias slice as fp32 in shared memory (one coalesced pass instead of 32 scalar
    // loads per thread and chunk), then walk the accumulator 32 columns at a time with the next chunk's tcgen05.ld already
    // in flight while the current one is converted and stored.
    const int q = warp & 3;                                 // this warp's TMEM lane quarter
    const int et = threadIdx.x - 64;                        // 0..127 among the epilogue threads
    int j = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x, ++j) {
      const int tile = u / p.splits;
      const int m0 = (tile % p.m_tiles) * BM, n0 = (tile / p.m_tiles) * p.bn;
      const int buf = j & 1;
      const int n_valid = min(p.bn, p.N - n0);
      const bool has_bias = p.bias != nullptr;
      if (has_bias) {
        named_bar_sync(1, 128);                             // everyone is done reading the previous tile's slice
        for (int i = et; i < p.bn; i += 128) {
          float bv = 0.f;
          if (i < n_valid)
            bv = p.bias_fp32 ? __ldg(static_cast<const float*>(p.bias) + n0 + i)
                             : __bfloat162float(static_cast<const __nv_bfloat16*>(p.bias)[n0 + i]);
          bias_s[i] = bv;
        }
        named_bar_sync(1, 128);
      }
      mbar_wait(&tmem_full_bar[buf], (j >> 1) & 1);
      tcgen05_fence_after();
      epilogue_tile(p, tmem_base + static_cast<uint32_t>(buf) * MAX_BN + (static_cast<uint32_t>(q * 32) << 16), m0 + q * 32 + lane, n0, n_valid,
                    has_bias, bias_s);
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]);
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------------------ CTA pairs
// The same GEMM with tcgen05.mma.cta_group::2: a cluster of two CTAs (one TPC) owns a 256 x bn tile. Each CTA stages ITS 128
// rows of A and HALF of B (bn/2 rows), the leader's elected thread issues M = 256 instructions that read both CTAs' shared
// memory, and each CTA's TMEM ends up with its own 128 rows of the accumulator. Per SM and K block that is 32 KB through the
// TMA / L2 path instead of 48 KB — the single-CTA kernel above is bound by exactly that traffic (tensor pipe 41-50 % active).
// Barriers: full[s] per CTA (TMA tx) -> a relay warp in each CTA forwards it to ready[s] on the leader (2 arrivals) ->
// issuer; empty[s] and tmem_full[b] are tcgen05.commit multicasts to both CTAs; tmem_empty[b] collects the 8 epilogue warps of
// the pair on the leader.
constexpr int NUM_THREADS_2 = 352;   // warp 0 producer, 1 issuer (leader) + TMEM owner, 2 relay, 3-10 epilogue (two per TMEM lane quarter:
                                     // with K = 768 a tile's main loop is as short as one warp's walk over its 8 column chunks)
constexpr int EPI_WARPS_2 = 8;
constexpr int STAGES_2 = 5;          // 5 x 32 KB ring + 8 warps x two 2 KB output boxes

template <bool A_MN, bool B_MN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS_2, 1)
gemm_bf16_2cta_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                      const __grid_constant__ CUtensorMap tmOut, const Params p) {
  pdl_launch_dependents();
  extern __shared__ __align__(1024) unsigned char smem_raw[];
  __shared__ __align__(8) uint64_t full_bar[STAGES_2], ready_bar[STAGES_2], empty_bar[STAGES_2], tmem_full_bar[2], tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) float bias_s[MAX_BN];

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = __shfl_sync(0xffffffffu, cluster_ctarank(), 0);
  const int pair = blockIdx.x >> 1, num_pairs = gridDim.x >> 1;
  const uint32_t ring = (smem_u32(smem_raw) + 1023u) & ~1023u;
  const int half_bn = p.bn >> 1;

  if (threadIdx.x == 0) {
    for (int s = 0; s < STAGES_2; ++s) { mbar_init(&full_bar[s], 1); mbar_init(&ready_bar[s], 2); mbar_init(&empty_bar[s], 1); }
    for (int b = 0; b < 2; ++b) { mbar_init(&tmem_full_bar[b], 1); mbar_init(&tmem_empty_bar[b], 2 * EPI_WARPS_2); }
    mbar_fence_init();
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.tma_store) tma_prefetch_desc(&tmOut);
  }
  if (warp == 1) {
    tmem_alloc_2cta(&tmem_base_slot, TMEM_COLS);
    tmem_relinquish_2cta();
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();   // the peer's barriers are initialised before anybody arrives on them remotely
  tcgen05_fence_after();
  pdl_wait();
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_slot, 0);
  const int total_units = p.total_units;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs: own A rows, own half of B)
    int s = 0;
    uint32_t ph = 0;
    for (int u = pair; u < total_units; u += num_pairs) {
      const int split = u % p.splits, tile = u / p.splits;
      const int m0 = (tile % p.m_tiles) * (2 * BM) + static_cast<int>(rank) * BM;
      const int nb0 = (tile / p.m_tiles) * p.bn + static_cast<int>(rank) * half_bn;
      const int kb0 = split * p.kb_per_split, kb1 = min(p.num_kb, kb0 + p.kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&empty_bar[s], ph ^ 1u);
        __syncwarp();
        if (elect_one()) {
          const uint32_t sa = ring + static_cast<uint32_t>(s) * p.stage_bytes, sb = sa + A_BYTES;
          mbar_expect_tx(&full_bar[s], p.stage_bytes);
          const int k0 = kb * BK;
          if (A_MN) {
#pragma unroll
            for (int b = 0; b < BM / 64; ++b)
              tma_load_2d(reinterpret_cast<void*>(__cvta_shared_to_generic(sa + b * BOX_BYTES)), &tmA, &full_bar[s], m0 + b * 64, k0);
          } else {
            tma_load_2d(reinterpret_cast<void*>(__cvta_shared_to_generic(sa)), &tmA, &full_bar[s], k0, m0);
          }
          if (B_MN) {
            for (int b = 0; b < half_bn / 64; ++b)
              tma_load_2d(reinterpret_cast<void*>(__cvta_shared_to_generic(sb + b * BOX_BYTES)), &tmB, &full_bar[s], nb0 + b * 64, k0);
          } else {
            tma_load_2d(reinterpret_cast<void*>(__cvta_shared_to_generic(sb)), &tmB, &full_bar[s], k0, nb0);
          }
        }
        __syncwarp();
        if (++s == STAGES_2) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only)
    if (rank == 0) {
      constexpr uint32_t a_lbo = A_MN ? BOX_BYTES : 16u, b_lbo = B_MN ? BOX_BYTES : 16u;
      constexpr uint32_t sbo = 1024u;
      constexpr uint32_t a_kstep = A_MN ? UMMA_K * 128u : UMMA_K * 2u, b_kstep = B_MN ? UMMA_K * 128u : UMMA_K * 2u;
      const uint32_t idesc = umma_idesc(UMMA_BF16, A_MN, B_MN, static_cast<uint32_t>(p.bn), 256);
      int s = 0;
      uint32_t ph = 0;
      int j = 0;
      for (int u = pair; u < total_units; u += num_pairs, ++j) {
        const int split = u % p.splits;
        const int kb0 = split * p.kb_per_split, kb1 = min(p.num_kb, kb0 + p.kb_per_split);
        const int buf = j & 1;
        mbar_wait(&tmem_empty_bar[buf], ((j >> 1) & 1) ^ 1u);
        tcgen05_fence_after();
        const uint32_t tmem_acc = tmem_base + static_cast<uint32_t>(buf) * MAX_BN;
        for (int kb = kb0; kb < kb1; ++kb) {
          mbar_wait(&ready_bar[s], ph);                            // both CTAs' tiles of this stage have landed
          tcgen05_fence_after();
          __syncwarp();
          const uint32_t sa = ring + static_cast<uint32_t>(s) * p.stage_bytes, sb = sa + A_BYTES;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; ++k) {
              const uint64_t da = umma_smem_desc(sa + k * a_kstep, a_lbo, sbo, UMMA_LAYOUT_SW128);
              const uint64_t db = umma_smem_desc(sb + k * b_kstep, b_lbo, sbo, UMMA_LAYOUT_SW128);
              umma_f16_ss_2cta(tmem_acc, da, db, idesc, (kb > kb0 || k > 0) ? 1u : 0u);
            }
            tcgen05_commit_2cta(&empty_bar[s], 0x3);               // both CTAs' slots are free once these MMAs have read them
            if (kb == kb1 - 1) tcgen05_commit_2cta(&tmem_full_bar[buf], 0x3);
          }
          __syncwarp();
          if (++s == STAGES_2) { s = 0; ph ^= 1u; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 2) {
    // ------------------------------------------------------------------ relay: this CTA's stage has landed -> tell the leader
    int s = 0;
    uint32_t ph = 0;
    for (int u = pair; u < total_units; u += num_pairs) {
      const int split = u % p.splits;
      const int kb0 = split * p.kb_per_split, kb1 = min(p.num_kb, kb0 + p.kb_per_split);
      for (int kb = kb0; kb < kb1; ++kb) {
        mbar_wait(&full_bar[s], ph);
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&ready_bar[s], 0);
        if (++s == STAGES_2) { s = 0; ph ^= 1u; }
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (both CTAs, own 128 rows)
    const int q = warp & 3;
    const int et = threadIdx.x - 96;                       // 0..255 among the epilogue threads
    const int half = (warp - 3) >> 2;                      // which of the two warps of this lane quarter
    int j = 0;
    for (int u = pair; u < total_units; u += num_pairs, ++j) {
      const int tile = u / p.splits;
      const int m0 = (tile % p.m_tiles) * (2 * BM) + static_cast<int>(rank) * BM, n0 = (tile / p.m_tiles) * p.bn;
      const int buf = j & 1;
      const int n_valid = min(p.bn, p.N - n0);
      const bool has_bias = p.bias != nullptr;
      if (has_bias) {
        named_bar_sync(1, 32 * EPI_WARPS_2);
        for (int i = et; i < p.bn; i += 32 * EPI_WARPS_2) {
          float bv = 0.f;
          if (i < n_valid)
            bv = p.bias_fp32 ? __ldg(static_cast<const float*>(p.bias) + n0 + i)
                             : __bfloat162float(static_cast<const __nv_bfloat16*>(p.bias)[n0 + i]);
          bias_s[i] = bv;
        }
        named_bar_sync(1, 32 * EPI_WARPS_2);
      }
      mbar_wait(&tmem_full_bar[buf], (j >> 1) & 1);
      tcgen05_fence_after();
      const uint32_t taddr = tmem_base + static_cast<uint32_t>(buf) * MAX_BN + (static_cast<uint32_t>(q * 32) << 16);
      if (p.tma_store)
        epilogue_tile_tma(p, &tmOut, taddr, m0 + q * 32, n0, n_valid, has_bias, bias_s,
                          smem_raw + (ring - smem_u32(smem_raw)) + static_cast<size_t>(STAGES_2) * p.stage_bytes + (warp - 3) * 4096, lane, half, 2);
      else
        epilogue_tile(p, taddr, m0 + q * 32 + lane, n0, n_valid, has_bias, bias_s, half, 2);
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive_cluster(&tmem_empty_bar[buf], 0);
    }
  }
  if (p.tma_store && warp >= 3 && lane == 0) tma_store_wait_all();   // this warp's output boxes have left shared memory
  // teardown: nobody may leave (or free TMEM) while the peer can still read this CTA's smem / arrive here
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_2cta(tmem_base, TMEM_COLS);
}

}  // namespace bf16gemm
}  // namespace rtdc

// ------------------------------------------------------------------------------------ host side
using namespace rtdc;
using namespace rtdc::bf16gemm;

namespace {

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                  const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = [] {
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess)
      return static_cast<EncodeTiledFn>(nullptr);
    return reinterpret_cast<EncodeTiledFn>(sym);
  }();
  return fn;
}

// [outer, inner] row-major bf16 matrix with row pitch ld elements; box = {box_inner, box_outer}, 128-byte swizzle.
int make_map_bf16(CUtensorMap* map, const void* base, int inner, int outer, int ld, int box_inner, int box_outer,
                  CUtensorMapSwizzle swizzle = CU_TENSOR_MAP_SWIZZLE_128B) {
  EncodeTiledFn enc = encode_fn();
  if (!enc) return fail(RTDC_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  if ((reinterpret_cast<uintptr_t>(base) & 15u) != 0) return fail(RTDC_ERR_ARG, "gemm_bf16: operand pointer must be 16-byte aligned");
  if ((static_cast<size_t>(ld) * 2) % 16 != 0) return fail(RTDC_ERR_ARG, "gemm_bf16: operand leading dimension must be a multiple of 8 elements (got %d)", ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 2};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(RTDC_ERR_CUDA, "cuTensorMapEncodeTiled (bf16) failed with CUresult %d", static_cast<int>(r));
  return RTDC_OK;
}

int sm_count() {
  static int n = [] {
    int dev = 0, v = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev);
    return v > 0 ? v : 148;
  }();
  return n;
}

template <bool A_MN, bool B_MN>
int launch2(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const Params& p, int pairs, size_t smem, cudaStream_t st) {
  static bool configured = false;
  if (!configured) {
    RTDC_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_2cta_kernel<A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         STAGES_2 * (A_BYTES + (MAX_BN / 2) * BK * 2) + EPI_WARPS_2 * 4096 + 1024));
    configured = true;
  }
  RTDC_CUDA_CHECK(launch_pdl(gemm_bf16_2cta_kernel<A_MN, B_MN>, dim3(2 * pairs), dim3(NUM_THREADS_2), smem, st, ta, tb, to, p));
  return RTDC_OK;
}

template <bool A_MN, bool B_MN>
int launch(const CUtensorMap& ta, const CUtensorMap& tb, const Params& p, int grid, size_t smem, cudaStream_t st) {
  static bool configured = false;
  if (!configured) {
    // the largest plan: four stages of a 128 x 64 A tile and a 256 x 64 B tile, plus the 1 KB alignment slack
    RTDC_CUDA_CHECK(cudaFuncSetAttribute(gemm_bf16_kernel<A_MN, B_MN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                         STAGES * (A_BYTES + MAX_BN * BK * 2) + 1024));
    configured = true;
  }
  RTDC_CUDA_CHECK(launch_pdl(gemm_bf16_kernel<A_MN, B_MN>, dim3(grid), dim3(NUM_THREADS), smem, st, ta, tb, p));
  return RTDC_OK;
}

}  // namespace

RTDC_API int rtdc_gemm_bf16(const rtdc_gemm_bf16_desc* d, void* stream) {
  if (!d || !d->a || !d->b || !d->out) return fail(RTDC_ERR_ARG, "gemm_bf16: null pointer");
  if (d->M <= 0 || d->N <= 0 || d->K <= 0) return fail(RTDC_ERR_ARG, "gemm_bf16: bad shape %d x %d x %d", d->M, d->N, d->K);
  if (d->act < 0 || d->act > 3) return fail(RTDC_ERR_ARG, "gemm_bf16: act must be 0 (none), 1 (relu), 2 (gelu tanh) or 3 (gelu erf)");
  Params p;
  memset(&p, 0, sizeof(p));
  p.M = d->M; p.N = d->N; p.K = d->K;
  // CTA pairs (256 x bn tiles, cta_group::2) whenever the problem has at least one such tile; RTDC_BF16_2CTA=0 keeps the
  // single-CTA kernel (128 x bn tiles), which also serves small M / N.
  static const bool allow_pair = [] { const char* e = getenv("RTDC_BF16_2CTA"); return !(e && e[0] == '0'); }();
  const bool pair = allow_pair && d->M >= 2 * BM && d->N >= 128;
  const int sms = sm_count();
  if (pair) {
    p.bn = d->N >= MAX_BN ? MAX_BN : round_up(d->N, 128);      // each CTA stages bn/2 rows of B: whole 64-row boxes when MN-major
    p.m_tiles = ceil_div(d->M, 2 * BM);
  } else {
    const int gran = d->b_mn ? 64 : 16;
    p.bn = d->N >= MAX_BN ? MAX_BN : round_up(d->N, gran);
    p.m_tiles = ceil_div(d->M, BM);
  }
  p.n_tiles = ceil_div(d->N, p.bn);
  p.num_kb = ceil_div(d->K, BK);
  const int tiles = p.m_tiles * p.n_tiles, slots = pair ? sms / 2 : sms;
  int splits = d->splits;
  if (splits <= 0) {
    splits = 1;
    // wgrad: few output tiles, long K (the batch's tokens): fill the machine once. (Measured and dropped: choosing the split that
    // minimises waves x K-blocks, e.g. 8 splits for 27 tiles — every extra split adds the whole tile once more with fp32
    // reductions, which cost more than the idle CTA pairs: 59.0 vs 51.5 us on the 2304 x 768 x 16384 wgrad.)
    if (d->out_fp32 && d->act == 0 && d->bias == nullptr && tiles * 2 <= slots) splits = slots / tiles;
  }
  if (splits > p.num_kb) splits = p.num_kb;
  if (splits > 1 && (!d->out_fp32 || d->act != 0 || d->bias != nullptr))
    return fail(RTDC_ERR_ARG, "gemm_bf16: a K split accumulates with fp32 atomics: it needs an fp32 output and no bias / activation");
  p.kb_per_split = ceil_div(p.num_kb, splits);
  p.splits = ceil_div(p.num_kb, p.kb_per_split);          // no empty split
  p.total_units = tiles * p.splits;
  const int b_rows = pair ? p.bn / 2 : p.bn;              // rows of B one CTA stages
  p.b_bytes = static_cast<uint32_t>(b_rows) * BK * 2;
  p.stage_bytes = A_BYTES + p.b_bytes;
  p.out = d->out; p.ldo = d->ldo; p.out_fp32 = d->out_fp32;
  p.bias = d->bias; p.bias_fp32 = d->bias_fp32;
  p.act = d->act;
  p.alpha = d->alpha == 0.f ? 1.f : d->alpha;
  if (d->ldo < d->N) return fail(RTDC_ERR_ARG, "gemm_bf16: ldo %d < N %d", d->ldo, d->N);
  CUtensorMap ta, tb;
  int rc = d->a_mn ? make_map_bf16(&ta, d->a, d->M, d->K, d->lda, 64, BK) : make_map_bf16(&ta, d->a, d->K, d->M, d->lda, BK, BM);
  if (rc != RTDC_OK) return rc;
  rc = d->b_mn ? make_map_bf16(&tb, d->b, d->N, d->K, d->ldb, 64, BK) : make_map_bf16(&tb, d->b, d->K, d->N, d->ldb, BK, b_rows);
  if (rc != RTDC_OK) return rc;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (p.splits > 1)   // the splits add into the output
    RTDC_CUDA_CHECK(cudaMemset2DAsync(d->out, static_cast<size_t>(d->ldo) * 4, 0, static_cast<size_t>(d->N) * 4, d->M, st));
  if (pair) {
    CUtensorMap to;
    memset(&to, 0, sizeof(to));
    static const bool allow_tma_store = [] { const char* e = getenv("RTDC_BF16_TMA_STORE"); return !(e && e[0] == '0'); }();
    if (allow_tma_store && !d->out_fp32 && p.splits == 1 && (reinterpret_cast<uintptr_t>(d->out) & 15u) == 0 && (static_cast<size_t>(d->ldo) * 2) % 16 == 0) {
      rc = make_map_bf16(&to, d->out, d->N, d->M, d->ldo, 32, 32, CU_TENSOR_MAP_SWIZZLE_64B);
      if (rc != RTDC_OK) return rc;
      p.tma_store = 1;
    }
    const size_t smem2 = static_cast<size_t>(STAGES_2) * p.stage_bytes + EPI_WARPS_2 * 4096 + 1024;
    const int pairs = p.total_units < slots ? p.total_units : slots;
    if (d->a_mn) return d->b_mn ? launch2<true, true>(ta, tb, to, p, pairs, smem2, st) : launch2<true, false>(ta, tb, to, p, pairs, smem2, st);
    return d->b_mn ? launch2<false, true>(ta, tb, to, p, pairs, smem2, st) : launch2<false, false>(ta, tb, to, p, pairs, smem2, st);
  }
  const size_t smem = static_cast<size_t>(STAGES) * p.stage_bytes + 1024;
  const int grid = p.total_units < sms ? p.total_units : sms;
  if (d->a_mn) return d->b_mn ? launch<true, true>(ta, tb, p, grid, smem, st) : launch<true, false>(ta, tb, p, grid, smem, st);
  return d->b_mn ? launch<false, true>(ta, tb, p, grid, smem, st) : launch<false, false>(ta, tb, p, grid, smem, st);
}
