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
#include <stdlib.h>

#include "common.h"
#include "ptx.cuh"
#include "../../include/rtdc_b200.h"

namespace rtdc {

constexpr int BM = 128;           // tile rows == TMEM lanes
// K block (fp32 elements per pipeline stage): 32 = one 128-byte swizzle row, 16 = one 64-byte row.
// 3xTF32 uses 16 so that four stages (raw + tail tiles) fit in shared memory.
constexpr int UMMA_K = 8;         // tf32: 32 bytes per instruction along K
constexpr int MAX_STAGES = 6;
constexpr int MAX_BN = 256;
constexpr int TMEM_COLS = 512;      // two 256-column accumulator buffers (epilogue of one tile overlaps the next main loop)
// 12 warps: 0 TMA producer, 1 MMA issuer, 2-7 splitters, 8-11 epilogue. Registers are allocated in groups of four warps, so
// a 10-warp CTA is charged for 12 (168 registers per thread either way): the two extra warps are free splitter throughput.
constexpr int NUM_THREADS = 384;
constexpr int SPLIT_THREADS = 192;
constexpr int HELP_THREADS = 128;   // splitter warps 2-5 (one per TMEM lane quarter) also take part of the final epilogue
constexpr int EPI_THREADS = 128;
constexpr int EPI_STAGING_BYTES = 32768;  // 4 epilogue warps x two 32x32 fp32 TMA-store boxes
constexpr int SMEM_LIMIT = 232448;  // 227 KB opt-in maximum per CTA
constexpr int STATIC_SMEM_RESERVE = 2048;  // barriers, bias slice, flags (static __shared__)

struct GemmParams {
  int M, N, K;
  int bn, stages, splits, kb_per_split, num_kb;
  int m_tiles, n_tiles;
  int m_tiles2;   // 256-row tiles (CTA-pair kernel)
  // weighted split-K: the tiles of the last (narrower) N column get their own split count, so that every work unit costs
  // about the same. Units are numbered: [tiles_full x splits] first, then [m_sched x splits_last].
  int m_sched, tiles_full, full_units, total_units, splits_last, kb_per_split_last;
  uint32_t a_bytes, b_bytes, stage_bytes;  // per stage: raw A, raw B, total (incl. tails when split)
  float* out; int ldo; int n_store;
  const float* bias; int relu;
  float drop_scale; uint32_t drop_thr16; uint32_t seed_lo, seed_hi; const uint32_t* step_ptr; uint32_t step_imm, salt;
  const float* mask_src; int ldm; float mask_scale;
  float* scratch;
  int debug;  // timing experiments only (RTDC_GEMM_DEBUG): 1 = no TMA loads, 2 = no MMAs, 4 = no global stores, 8 = no split pass, 16 = no proxy fence
  float* colsum_partial; int ld_colsum;   // optional: per (m_tile, warp) column sums of the output tile rows, [4*m_tiles][ld]
  // exact operands (no tail tile, one MMA fewer per K step) and the affine map of the accumulator
  int a_exact, b_exact;
  uint32_t a_tail_off, b_tail_off;        // tail tiles, bytes from the start of a stage
  uint32_t split_lo_vec, split_hi_vec;    // float4 range of the raw tiles the splitters have to process
  float alpha; const float* row_bias; float row_bias_scale;
  // MN-major operands: tiles that lie inside the first a3_limit / b3_limit rows (whole 32-element blocks) are loaded by
  // ONE 3-D TMA instruction ([blocks][k][32]) instead of one 2-D box per 32 rows; 0 = no 3-D map
  int a3_limit, b3_limit;
  uint32_t* mask_bits_out; int ld_mask_out;        // bit-packed (out > 0), one word per 32 columns
  const uint32_t* mask_bits_in; int ld_mask_in;    // bit-packed backward mask (instead of mask_src)
  int tma_store;         // epilogue goes through smem + cp.async.bulk.tensor stores (tmOut valid)
  long long* dbg_times;  // RTDC_GEMM_DEBUG_BUF: per-CTA clock64 stamps [grid][8]
};

struct Unit { int tile, split, kb_begin, n_iters; };
// Work unit -> (tile, K split, K-block range). Tiles are numbered n_tile * m_sched + m_idx, so the last N column's tiles
// are the last m_sched ones.
__device__ __forceinline__ Unit decode_unit(const GemmParams& p, int u) {
  Unit w;
  int kps;
  if (u < p.full_units) {
    w.tile = u % p.tiles_full; w.split = u / p.tiles_full; kps = p.kb_per_split;
  } else {
    const int v = u - p.full_units;
    w.tile = p.tiles_full + v % p.m_sched; w.split = v / p.m_sched; kps = p.kb_per_split_last;
  }
  w.kb_begin = w.split * kps;
  w.n_iters = min(p.num_kb, w.kb_begin + kps) - w.kb_begin;
  return w;
}

// keep flag for dropout: 16 random bits per element, 8 elements per Philox call.
__device__ __forceinline__ uint32_t rand16(const Philox4& r, int j) {
  const uint32_t w = (j < 2) ? r.x : (j < 4) ? r.y : (j < 6) ? r.z : r.w;
  return (j & 1) ? (w >> 16) : (w & 0xFFFFu);
}

// Dropout keep flags of 32 consecutive elements of one row (bit q = keep column col0 + q). Element index e = row*N + col,
// 8 elements per Philox call, 16 bits each.
__device__ __forceinline__ uint32_t dropout_keep_bits(const GemmParams& p, int row, int col0, uint32_t step) {
  const unsigned long long e_base = static_cast<unsigned long long>(row) * static_cast<unsigned long long>(p.N) + col0;
  uint32_t bits = 0;
  if ((e_base & 7ull) == 0ull) {
#pragma unroll
    for (int g = 0; g < 4; ++g) {
      const unsigned long long grp = (e_base >> 3) + g;
      const Philox4 rnd = philox4x32_10(static_cast<uint32_t>(grp), static_cast<uint32_t>(grp >> 32), step, p.salt, p.seed_lo, p.seed_hi);
#pragma unroll
      for (int q = 0; q < 8; ++q) bits |= (rand16(rnd, q) >= p.drop_thr16 ? 1u : 0u) << (g * 8 + q);
    }
  } else {
    // unaligned row start (N not a multiple of 8): one Philox call per element (rare shapes only)
#pragma unroll 1
    for (int q = 0; q < 32; ++q) {
      const unsigned long long e = e_base + q, grp = e >> 3;
      const Philox4 rnd = philox4x32_10(static_cast<uint32_t>(grp), static_cast<uint32_t>(grp >> 32), step, p.salt, p.seed_lo, p.seed_hi);
      bits |= (rand16(rnd, static_cast<int>(e & 7ull)) >= p.drop_thr16 ? 1u : 0u) << q;
    }
  }
  return bits;
}
// The keep flags of a whole unit (up to eight 32-column chunks), computed by the epilogue warps while they wait for the
// accumulators, so that the Philox rounds are off the critical path between the last MMA and the last store.
struct KeepBits { uint32_t w[8]; };
__device__ __forceinline__ uint32_t keep_word(const KeepBits& k, int c) {
  uint32_t r = k.w[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) r = (c == i) ? k.w[i] : r;
  return r;
}
__device__ __forceinline__ void precompute_keep_bits(const GemmParams& p, int row, int n0, uint32_t step, KeepBits& k) {
  const int n_chunks = (min(p.bn, p.n_store - n0) + 31) / 32;
#pragma unroll
  for (int c = 0; c < 8; ++c) k.w[c] = (c < n_chunks && row < p.M) ? dropout_keep_bits(p, row, n0 + c * 32, step) : 0u;
}

// Backward-mask flags of a whole unit (bit q of word c = mask_src[row, n0 + 32c + q] > 0), also gathered while the epilogue
// warps wait for the accumulators: the loads then have the whole main loop to land instead of one chunk's time each
// (one chunk in flight at ~1 us of memory latency made the dgrad epilogue 14 us per unit against 6 us without a mask).
__device__ __forceinline__ void precompute_mask_bits(const GemmParams& p, int row, int n0, KeepBits& k) {
  const int n_chunks = (min(p.bn, p.n_store - n0) + 31) / 32;
  const bool vec_ok = ((reinterpret_cast<uintptr_t>(p.mask_src) & 15u) == 0) && (p.ldm & 3) == 0 && (n0 & 3) == 0;
  const float* ms = p.mask_src + static_cast<size_t>(row < p.M ? row : 0) * p.ldm + n0;
#pragma unroll
  for (int c = 0; c < 8; ++c) {
    uint32_t bits = 0;
    if (c < n_chunks && row < p.M) {
      if (vec_ok && n0 + c * 32 + 32 <= p.N) {
        float4 m[8];
#pragma unroll
        for (int q = 0; q < 8; ++q) m[q] = __ldg(reinterpret_cast<const float4*>(ms + c * 32) + q);
#pragma unroll
        for (int q = 0; q < 8; ++q)
          bits |= (m[q].x > 0.f ? 1u : 0u) << (4 * q) | (m[q].y > 0.f ? 2u : 0u) << (4 * q) | (m[q].z > 0.f ? 4u : 0u) << (4 * q) |
                  (m[q].w > 0.f ? 8u : 0u) << (4 * q);
      } else {
#pragma unroll 1
        for (int q = 0; q < 32; ++q)
          if (n0 + c * 32 + q < p.N && __ldg(ms + c * 32 + q) > 0.f) bits |= 1u << q;
      }
    }
    k.w[c] = bits;
  }
}

// Bit-packed backward mask: this row's words for the unit's chunks (32 bytes instead of 1 KB of activations per thread).
__device__ __forceinline__ void load_mask_words(const GemmParams& p, int row, int n0, KeepBits& k) {
  const int n_chunks = (min(p.bn, p.n_store - n0) + 31) / 32;
  const uint32_t* src = p.mask_bits_in + static_cast<size_t>(row < p.M ? row : 0) * p.ld_mask_in + (n0 >> 5);
#pragma unroll
  for (int c = 0; c < 8; ++c) k.w[c] = (c < n_chunks && row < p.M && n0 + c * 32 < p.N) ? __ldg(src + c) : 0u;
}

// Epilogue of one work unit, run by the four epilogue warps: TMEM -> registers -> (+bias, ReLU, dropout,
// backward mask) -> swizzled smem box -> TMA store (or split-K partial to scratch). `release()` hands the
// accumulator buffer back to the MMA issuer right after this warp's last TMEM read. `part` selects the 32-column
// chunks this call covers: 0 = all, 1 = first half, 2 = second half (the CTA's final unit is shared with the splitter
// warps, which have nothing left to split by then).
template <class ReleaseFn>
__device__ __forceinline__ void epilogue_unit(const GemmParams& p, const CUtensorMap* tmOut, uint32_t taddr, int umma_n,
                                              int tile, int split, int m0, int n0, int r_in_tile, int lane, int wq,
                                              const float* s_bias, uint8_t* stage_buf, uint32_t& stores_issued,
                                              uint32_t step, ReleaseFn release, int part = 0, const KeepBits* keep = nullptr,
                                              const KeepBits* mask_bits = nullptr) {
  const bool has_drop = p.drop_thr16 != 0u;
  const bool has_mask = p.mask_src != nullptr || p.mask_bits_in != nullptr;
  const bool relu = p.relu != 0;
  const int row = m0 + r_in_tile;
  const int n_cols_out = min(p.bn, p.n_store - n0);  // columns to write (valid + zero padding)
  const int n_chunks = (n_cols_out + 31) / 32;
  const int acc_chunks = (umma_n + 31) / 32;         // chunks that hold MMA output
  const float alpha = p.alpha;
  const float rb = (p.row_bias != nullptr && row < p.M) ? __ldg(p.row_bias + row) * p.row_bias_scale : 0.f;
  uint32_t r[32];
  if (p.splits > 1) {
    // split-K: park this split's partial tile in scratch; splitk_reduce_kernel sums the splits in
    // order (deterministic) and applies the epilogue
    float* my = p.scratch + (static_cast<size_t>(tile < 0 ? 0 : tile) * p.splits + split) * (BM * MAX_BN);
    const int total = tile < 0 ? 0 : acc_chunks, half = (total + 1) >> 1;
    for (int c = (part == 2 ? half : 0); c < (part == 1 ? half : total); ++c) {
      tmem_ld_32x32(taddr + c * 32, r);
      tmem_ld_wait();
      float4* dst = reinterpret_cast<float4*>(my + (static_cast<size_t>(c) * BM + r_in_tile) * 32);
#pragma unroll
      for (int q = 0; q < 8; ++q)
        dst[q] = make_float4(__uint_as_float(r[4 * q]), __uint_as_float(r[4 * q + 1]),
                             __uint_as_float(r[4 * q + 2]), __uint_as_float(r[4 * q + 3]));
    }
    tcgen05_fence_before();
    __syncwarp();
    release();
    return;
  }

  // backward mask of chunk 0 (the loads of chunk c+1 are issued while chunk c is being processed)
  const bool mask_vec_ok = p.mask_src != nullptr && row < p.M && ((reinterpret_cast<uintptr_t>(p.mask_src) & 15u) == 0) && (p.ldm & 3) == 0 &&
                           (n0 & 3) == 0;
  // the epilogue warps arrive with their dropout flags precomputed, the helping splitter warps do not: 5/8 vs 3/8
  const int c_half = (has_drop || p.mask_src != nullptr) ? (5 * n_chunks + 7) >> 3 : (n_chunks + 1) >> 1;
  const int c_begin = part == 2 ? c_half : 0, c_end = part == 1 ? c_half : n_chunks;
  float4 mk[8];
  if (mask_bits == nullptr && mask_vec_ok && c_begin < c_end && n0 + c_begin * 32 + 32 <= p.N) {
    const float4* ms4 = reinterpret_cast<const float4*>(p.mask_src + static_cast<size_t>(row) * p.ldm + n0 + c_begin * 32);
#pragma unroll
    for (int q = 0; q < 8; ++q) mk[q] = __ldg(ms4 + q);
  }
  for (int c = c_begin; c < c_end; ++c) {
    float v[32];
    if (c < acc_chunks) {
      tmem_ld_32x32(taddr + c * 32, r);
      tmem_ld_wait();
    }
    if (c == acc_chunks - 1 || (acc_chunks == 0 && c == 0)) {
      // last TMEM read of this unit: hand the accumulator buffer back to the MMA warp
      tcgen05_fence_before();
      __syncwarp();
      release();
    }
    // TMEM columns past the MMA's N were never written: force them (and their bias) to zero.
    // Columns between N and umma_n hold exact zeros already (TMA zero-fills out-of-range B rows).
    const int live = umma_n - c * 32;
    if (live >= 32) {
      // whole chunk live: bias through eight 128-bit shared loads (independent of one another; the element-wise
      // form below compiles to a serial chain of predicated scalar loads)
      float4 bb[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) bb[q] = reinterpret_cast<const float4*>(s_bias + c * 32)[q];
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const float x0 = fmaf(__uint_as_float(r[4 * q]), alpha, bb[q].x) + rb;
        const float x1 = fmaf(__uint_as_float(r[4 * q + 1]), alpha, bb[q].y) + rb;
        const float x2 = fmaf(__uint_as_float(r[4 * q + 2]), alpha, bb[q].z) + rb;
        const float x3 = fmaf(__uint_as_float(r[4 * q + 3]), alpha, bb[q].w) + rb;
        v[4 * q] = relu ? fmaxf(x0, 0.f) : x0;
        v[4 * q + 1] = relu ? fmaxf(x1, 0.f) : x1;
        v[4 * q + 2] = relu ? fmaxf(x2, 0.f) : x2;
        v[4 * q + 3] = relu ? fmaxf(x3, 0.f) : x3;
      }
    } else {
#pragma unroll
      for (int q = 0; q < 32; ++q) v[q] = (c < acc_chunks) ? __uint_as_float(r[q]) : 0.f;
#pragma unroll
      for (int q = 0; q < 32; ++q) {
        float x = (q < live) ? fmaf(v[q], alpha, s_bias[c * 32 + q]) + rb : 0.f;
        v[q] = relu ? fmaxf(x, 0.f) : x;
      }
    }
    const int col0 = n0 + c * 32;
    // dropout (flags precomputed by the caller while it waited for the accumulators, or computed here)
    if (has_drop && row < p.M) {
      const uint32_t bits = keep != nullptr ? keep_word(*keep, c) : dropout_keep_bits(p, row, col0, step);
#pragma unroll
      for (int q = 0; q < 32; ++q) v[q] = ((bits >> q) & 1u) ? v[q] * p.drop_scale : 0.f;
    }
    // backward mask of the previous layer's ReLU+dropout output
    if (has_mask && mask_bits != nullptr) {
      const uint32_t bits = keep_word(*mask_bits, c);
#pragma unroll
      for (int q = 0; q < 32; ++q) v[q] = ((bits >> q) & 1u) ? v[q] * p.mask_scale : 0.f;
    } else if (has_mask && row < p.M) {
      const float* ms = p.mask_src + static_cast<size_t>(row) * p.ldm + col0;
      if (mask_vec_ok && col0 + 32 <= p.N) {
#pragma unroll
        for (int q = 0; q < 8; ++q) {
          v[4 * q] = mk[q].x > 0.f ? v[4 * q] * p.mask_scale : 0.f;
          v[4 * q + 1] = mk[q].y > 0.f ? v[4 * q + 1] * p.mask_scale : 0.f;
          v[4 * q + 2] = mk[q].z > 0.f ? v[4 * q + 2] * p.mask_scale : 0.f;
          v[4 * q + 3] = mk[q].w > 0.f ? v[4 * q + 3] * p.mask_scale : 0.f;
        }
      } else {
#pragma unroll
        for (int q = 0; q < 32; ++q)
          if (col0 + q < p.N) v[q] = (__ldg(ms + q) > 0.f) ? v[q] * p.mask_scale : 0.f;
      }
      // next chunk's mask: in flight during this chunk's stores and the next TMEM load
      if (mask_vec_ok && c + 1 < c_end && col0 + 64 <= p.N) {
#pragma unroll
        for (int q = 0; q < 8; ++q) mk[q] = __ldg(reinterpret_cast<const float4*>(ms + 32) + q);
      }
    }
    // bit-packed (out > 0) for a later dgrad
    if (p.mask_bits_out != nullptr && row < p.M && col0 < p.N) {
      uint32_t bits = 0;
#pragma unroll
      for (int q = 0; q < 32; ++q) bits |= (v[q] > 0.f ? 1u : 0u) << q;   // columns past N hold zeros
      p.mask_bits_out[static_cast<size_t>(row) * p.ld_mask_out + (col0 >> 5)] = bits;
    }
    if (p.debug & 4) continue;
    if (p.tma_store) {
      // registers -> 128B-swizzled smem box -> one bulk tensor store per warp (clipped at the
      // tensor edge by the TMA unit, full 128-byte lines on the wire)
      uint8_t* box = stage_buf + (stores_issued & 1u) * 4096;
      if (stores_issued >= 2) {
        if (lane == 0) tma_store_wait_read<1>();  // the store issued two boxes ago has drained this one
        __syncwarp();
      }
      uint8_t* rowp = box + lane * 128;
#pragma unroll
      for (int q = 0; q < 8; ++q)
        *reinterpret_cast<float4*>(rowp + ((q ^ (lane & 7)) << 4)) = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
      fence_proxy_async_smem();
      __syncwarp();
      if (lane == 0) {
        tma_store_2d(tmOut, box, col0, m0 + wq * 32);
        tma_store_commit();
      }
      ++stores_issued;
      if (p.colsum_partial != nullptr && m0 < p.M) {   // (a CTA pair's second half can lie entirely past M)
        // column sums of this warp's 32 x 32 box straight from the staged tile (lane = column; the swizzle
        // makes the 32 lanes hit 32 different banks): the bias gradient costs no extra pass over HBM
        const int valid = min(32, p.M - (m0 + wq * 32));
        // four independent partial sums (a single chain of 32 dependent shared loads + adds costs ~1000 cycles per chunk)
        float c4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int rr = 0; rr < 32; ++rr) {
          const float e = *reinterpret_cast<const float*>(box + rr * 128 + ((((lane >> 2) ^ (rr & 7))) << 4) + ((lane & 3) << 2));
          c4[rr & 3] += rr < valid ? e : 0.f;     // rows past M hold bias-only values that the store clips
        }
        const float cs = (c4[0] + c4[1]) + (c4[2] + c4[3]);
        if (col0 + lane < p.N)
          p.colsum_partial[static_cast<size_t>((m0 / BM) * 4 + wq) * p.ld_colsum + col0 + lane] = cs;
      }
    } else if (row < p.M) {
      float* dst = p.out + static_cast<size_t>(row) * p.ldo + col0;
      const int rem = n_cols_out - c * 32;
      if (rem >= 32 && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0)) {
#pragma unroll
        for (int q = 0; q < 8; ++q)
          reinterpret_cast<float4*>(dst)[q] = make_float4(v[4 * q], v[4 * q + 1], v[4 * q + 2], v[4 * q + 3]);
      } else {
#pragma unroll
        for (int q = 0; q < 32; ++q)
          if (q < rem) dst[q] = v[q];
      }
    }
  }
}

template <bool A_MN, bool B_MN, int NSPLIT, int BK>
__global__ void __launch_bounds__(NUM_THREADS, 1)
gemm_tf32_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmA3,
                 const __grid_constant__ CUtensorMap tmB3, const GemmParams p) {
  extern __shared__ uint8_t smem_dyn[];
  pdl_launch_dependents();
  __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t ready_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];    // MMA -> epilogue: accumulator buffer complete
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];   // epilogue -> MMA: accumulator buffer drained
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) float s_bias[MAX_BN];

  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);   // warp-uniform for the compiler
  const int lane = threadIdx.x & 31;

  // 1024-byte aligned stage ring (swizzle atoms are up to 1024 B), epilogue staging right behind it
  const uint32_t dyn_base = smem_u32(smem_dyn);
  const uint32_t ring = (dyn_base + 1023u) & ~1023u;
  uint8_t* ring_ptr = smem_dyn + (ring - dyn_base);
  uint8_t* staging = ring_ptr + static_cast<size_t>(p.stages) * p.stage_bytes;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&ready_bar[s], SPLIT_THREADS / 32);
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], EPI_THREADS / 32);
    }
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.tma_store) tma_prefetch_desc(&tmOut);
    if (p.a3_limit) tma_prefetch_desc(&tmA3);
    if (p.b3_limit) tma_prefetch_desc(&tmB3);
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
  long long* stamps = p.dbg_times ? p.dbg_times + static_cast<size_t>(blockIdx.x) * 8 : nullptr;
  if (stamps && threadIdx.x == 0) { stamps[0] = clock64(); stamps[7] = static_cast<long long>(global_timer_ns()); }

  // A CTA walks over work units (output tile x K split) round-robin; every role derives the same
  // sequence, so the smem ring and the two TMEM accumulator buffers just keep cycling across units and
  // the epilogue of unit j overlaps the main loop of unit j+1.
  const int total_units = p.total_units;
  // the splitter warps help with the CTA's final epilogue when the ring can host their four store boxes
  const bool share_final = !(p.debug & 32) && static_cast<uint32_t>(p.stages) * p.stage_bytes >= 32768u;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    // (the whole warp walks the loop, one elected lane issues: operands stay in uniform registers)
    {
      int s = 0;
      uint32_t ph = 0;
      for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
        const Unit w = decode_unit(p, u);
        const int tile = w.tile, kb_begin = w.kb_begin, n_iters = w.n_iters;
        const int m0 = (tile % p.m_tiles) * BM, n0 = (tile / p.m_tiles) * p.bn;
        for (int it = 0; it < n_iters; ++it) {
          mbar_wait(&empty_bar[s], ph ^ 1u);
          __syncwarp();
          uint8_t* sa = ring_ptr + static_cast<size_t>(s) * p.stage_bytes;
          uint8_t* sb = sa + p.a_bytes;
          if (!elect_one()) {
          } else if (p.debug & 1) {
            mbar_arrive(&full_bar[s]);
          } else {
            mbar_expect_tx(&full_bar[s], p.a_bytes + p.b_bytes);
            const int k0 = (kb_begin + it) * BK;
            if (!A_MN) {
              tma_load_2d(sa, &tmA, &full_bar[s], k0, m0);
            } else if (m0 + BM <= p.a3_limit) {
              tma_load_3d(sa, &tmA3, &full_bar[s], 0, k0, m0 >> 5);
            } else {
#pragma unroll
              for (int b = 0; b < BM / 32; ++b) tma_load_2d(sa + b * (128 * BK), &tmA, &full_bar[s], m0 + 32 * b, k0);
            }
            if (!B_MN) {
              tma_load_2d(sb, &tmB, &full_bar[s], k0, n0);
            } else if (n0 + p.bn <= p.b3_limit) {
              tma_load_3d(sb, &tmB3, &full_bar[s], 0, k0, n0 >> 5);
            } else {
              for (int b = 0; b < p.bn / 32; ++b) tma_load_2d(sb + b * (128 * BK), &tmB, &full_bar[s], n0 + 32 * b, k0);
            }
          }
          __syncwarp();
          if (++s == p.stages) { s = 0; ph ^= 1u; }
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    // (converged warp + elect.sync: no per-MMA R2UR election loop, see the CTA-pair kernel)
    {
      const bool a_exact = p.a_exact != 0, b_exact = p.b_exact != 0;
      // K-major: 8-row groups (SW128 for 128-byte rows, SW64 for 64-byte rows). MN-major: 32-wide MN blocks
      // one TMA box apart, 4-row K groups 512 B apart (SW128 with 32-byte atoms: the only MN-major layout
      // tf32 accepts).
      constexpr uint32_t kmajor_sbo = 8u * BK * 4u;
      constexpr uint32_t kmajor_lt = BK == 32 ? UMMA_LAYOUT_SW128 : UMMA_LAYOUT_SW64;
      constexpr uint32_t mn_box = 128u * BK;
      const uint32_t a_lbo = A_MN ? mn_box : 16u, b_lbo = B_MN ? mn_box : 16u;
      const uint32_t a_sbo = A_MN ? 512u : kmajor_sbo, b_sbo = B_MN ? 512u : kmajor_sbo;
      const uint32_t a_lt = A_MN ? UMMA_LAYOUT_SW128_BASE32B : kmajor_lt;
      const uint32_t b_lt = B_MN ? UMMA_LAYOUT_SW128_BASE32B : kmajor_lt;
      const uint32_t a_kstep = A_MN ? 1024u : 32u, b_kstep = B_MN ? 1024u : 32u;
      int s = 0;
      uint32_t ph = 0;
      int j = 0;
      for (int u = blockIdx.x; u < total_units; u += gridDim.x, ++j) {
        const Unit w = decode_unit(p, u);
        const int tile = w.tile, kb_begin = w.kb_begin, n_iters = w.n_iters;
        const int n0 = (tile / p.m_tiles) * p.bn;
        const int n_valid = min(p.bn, p.N - n0);
        const uint32_t umma_n = static_cast<uint32_t>(min(p.bn, (n_valid + 15) & ~15));
        const uint32_t idesc = umma_idesc(UMMA_TF32, A_MN, B_MN, umma_n);
        const int buf = j & 1;
        mbar_wait(&tmem_empty_bar[buf], ((j >> 1) & 1) ^ 1u);   // the epilogue has drained this buffer
        tcgen05_fence_after();
        const uint32_t tmem_acc = tmem_base + static_cast<uint32_t>(buf) * MAX_BN;
        for (int it = 0; it < n_iters; ++it) {
          mbar_wait(NSPLIT == 3 ? &ready_bar[s] : &full_bar[s], ph);
          tcgen05_fence_after();
          __syncwarp();
          const uint32_t sa = ring + static_cast<uint32_t>(s) * p.stage_bytes;
          const uint32_t sb = sa + p.a_bytes;
          if (!elect_one()) {
          } else if (p.debug & 2) {
            mbar_arrive(&empty_bar[s]);
          } else {
            const int k_left = p.K - (kb_begin + it) * BK;
            const int ksteps = min(BK / UMMA_K, (k_left + UMMA_K - 1) / UMMA_K);
            for (int k = 0; k < ksteps; ++k) {
              const uint64_t da = umma_smem_desc(sa + k * a_kstep, a_lbo, a_sbo, a_lt);
              const uint64_t db = umma_smem_desc(sb + k * b_kstep, b_lbo, b_sbo, b_lt);
              const uint32_t acc = (it > 0 || k > 0) ? 1u : 0u;
              if (NSPLIT == 3) {
                // small terms first; an exact operand has no tail, so its correction MMA is dropped
                uint32_t acc2 = acc;
                if (!a_exact) {
                  const uint64_t da_t = umma_smem_desc(sa + p.a_tail_off + k * a_kstep, a_lbo, a_sbo, a_lt);
                  umma_tf32_ss(tmem_acc, da_t, db, idesc, acc2);
                  acc2 = 1u;
                }
                if (!b_exact) {
                  const uint64_t db_t = umma_smem_desc(sa + p.b_tail_off + k * b_kstep, b_lbo, b_sbo, b_lt);
                  umma_tf32_ss(tmem_acc, da, db_t, idesc, acc2);
                  acc2 = 1u;
                }
                umma_tf32_ss(tmem_acc, da, db, idesc, acc2);
              } else {
                umma_tf32_ss(tmem_acc, da, db, idesc, acc);
              }
            }
            tcgen05_commit(&empty_bar[s]);  // smem slot reusable once these MMAs have read it
          }
          __syncwarp();
          if (++s == p.stages) { s = 0; ph ^= 1u; }
        }
        if (elect_one()) {
          if (p.debug & 2) mbar_arrive(&tmem_full_bar[buf]);
          else tcgen05_commit(&tmem_full_bar[buf]);  // accumulators of this unit complete
          if (stamps && j == 0) stamps[2] = clock64();
        }
        __syncwarp();
      }
    }
    __syncwarp();
  } else if (warp < 2 + SPLIT_THREADS / 32) {
    // ------------------------------------------------------------------ splitters (3xTF32 only)
    if (NSPLIT == 3) {
      const int st = threadIdx.x - 64;  // 0..127
      const uint32_t raw_bytes = p.a_bytes + p.b_bytes;
      // raw float4 range to split ([A|B], minus an exact operand); tail[i] pairs with raw[i] for both layouts
      const int lo = static_cast<int>(p.split_lo_vec), nvec = static_cast<int>(p.split_hi_vec);
      int s = 0;
      uint32_t ph = 0;
      for (int u = blockIdx.x; u < total_units; u += gridDim.x) {
        const int n_iters = decode_unit(p, u).n_iters;
        for (int it = 0; it < n_iters; ++it) {
          mbar_wait(&full_bar[s], ph);
          const float4* raw = reinterpret_cast<const float4*>(ring_ptr + static_cast<size_t>(s) * p.stage_bytes);
          float4* tail = reinterpret_cast<float4*>(ring_ptr + static_cast<size_t>(s) * p.stage_bytes + raw_bytes) - lo;
          // SPLIT_ILP loads are issued back to back before any store: the compiler cannot prove that the
          // tail tile does not alias the raw tile, so a plain loop serialises LDS -> STS -> LDS.
          constexpr int SPLIT_ILP = 6;
          for (int base = lo + st; base < ((p.debug & 8) ? 0 : nvec); base += SPLIT_THREADS * SPLIT_ILP) {
            float4 v[SPLIT_ILP];
#pragma unroll
            for (int q = 0; q < SPLIT_ILP; ++q) {
              const int i = base + q * SPLIT_THREADS;
              if (i < nvec) v[q] = raw[i];
            }
#pragma unroll
            for (int q = 0; q < SPLIT_ILP; ++q) {
              const int i = base + q * SPLIT_THREADS;
              if (i < nvec) {
                float4 t;
                t.x = v[q].x - __uint_as_float(__float_as_uint(v[q].x) & 0xFFFFE000u);
                t.y = v[q].y - __uint_as_float(__float_as_uint(v[q].y) & 0xFFFFE000u);
                t.z = v[q].z - __uint_as_float(__float_as_uint(v[q].z) & 0xFFFFE000u);
                t.w = v[q].w - __uint_as_float(__float_as_uint(v[q].w) & 0xFFFFE000u);
                tail[i] = t;
              }
            }
          }
          fence_proxy_async_smem();  // generic-proxy writes -> visible to the tensor core's async proxy
          __syncwarp();
          if (lane == 0) mbar_arrive(&ready_bar[s]);
          if (++s == p.stages) { s = 0; ph ^= 1u; }
        }
      }
    }
    // nothing left to split: take the second half of this CTA's final epilogue (staging boxes in the idle ring)
    if (share_final && warp < 2 + HELP_THREADS / 32) {
      const int n_mine = (total_units - static_cast<int>(blockIdx.x) + static_cast<int>(gridDim.x) - 1) / static_cast<int>(gridDim.x);
      const int j = n_mine - 1, u = static_cast<int>(blockIdx.x) + j * static_cast<int>(gridDim.x);
      const Unit w = decode_unit(p, u);
      const int tile = w.tile, split = w.split;
      const int m0 = (tile % p.m_tiles) * BM, n0 = (tile / p.m_tiles) * p.bn;
      const int n_valid = min(p.bn, p.N - n0);
      const int umma_n = min(p.bn, (n_valid + 15) & ~15);
      const int buf = j & 1;
      const int wq = warp & 3;
      const uint32_t taddr = tmem_base + static_cast<uint32_t>(buf) * MAX_BN + (static_cast<uint32_t>(wq * 32) << 16);
      uint32_t step = p.step_imm;
      if (p.drop_thr16 != 0u && p.step_ptr != nullptr) step = *p.step_ptr;
      uint32_t stores_issued = 0;
      KeepBits hmb;
      if (p.mask_bits_in != nullptr) load_mask_words(p, m0 + wq * 32 + lane, n0, hmb);
      named_bar_sync(2, HELP_THREADS + EPI_THREADS);   // the epilogue warps have staged this unit's bias slice
      mbar_wait(&tmem_full_bar[buf], (j >> 1) & 1);     // all MMAs done: accumulators complete, ring idle
      tcgen05_fence_after();
      epilogue_unit(p, &tmOut, taddr, umma_n, tile, split, m0, n0, wq * 32 + lane, lane, wq, s_bias, ring_ptr + static_cast<size_t>(wq) * 8192,
                    stores_issued, step, []() {}, 2, nullptr, p.mask_bits_in != nullptr ? &hmb : nullptr);
      if (p.tma_store && lane == 0) tma_store_wait_all();
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps
    const int et = threadIdx.x - (64 + SPLIT_THREADS);  // 0..127
    const int wq = warp & 3;                            // TMEM lane quarter this warp may read
    const int r_in_tile = wq * 32 + lane;
    uint32_t step = p.step_imm;
    if (p.drop_thr16 != 0u && p.step_ptr != nullptr) step = *p.step_ptr;
    uint8_t* stage_buf = staging + static_cast<size_t>(wq) * 8192;   // two 32x32 fp32 boxes per warp
    uint32_t stores_issued = 0;
    int j = 0;
    for (int u = blockIdx.x; u < total_units; u += gridDim.x, ++j) {
      const Unit w = decode_unit(p, u);
      const int tile = w.tile, split = w.split;
      const int m0 = (tile % p.m_tiles) * BM, n0 = (tile / p.m_tiles) * p.bn;
      const int n_valid = min(p.bn, p.N - n0);
      const int umma_n = min(p.bn, (n_valid + 15) & ~15);
      const int buf = j & 1;
      const uint32_t taddr = tmem_base + static_cast<uint32_t>(buf) * MAX_BN + (static_cast<uint32_t>(wq * 32) << 16);

      // bias slice of this tile -> smem (zero beyond N); the barrier also fences the previous unit's reads
      const bool shared = share_final && u + static_cast<int>(gridDim.x) >= total_units;   // final unit: split with the splitter warps
      named_bar_sync(1, EPI_THREADS);
      for (int c = et; c < MAX_BN; c += EPI_THREADS)
        s_bias[c] = (p.bias != nullptr && c < p.bn && n0 + c < p.N) ? __ldg(p.bias + n0 + c) : 0.f;
      // dropout flags first (idle time), then the barrier: on the final unit it also waits for the splitter warps
      KeepBits kb;
      const bool pre = p.drop_thr16 != 0u && p.splits == 1;
      if (pre) precompute_keep_bits(p, m0 + r_in_tile, n0, step, kb);
      KeepBits mb;
      const bool pre_mask = (p.mask_src != nullptr || p.mask_bits_in != nullptr) && p.splits == 1;
      if (p.mask_bits_in != nullptr) load_mask_words(p, m0 + r_in_tile, n0, mb);
      else if (pre_mask) precompute_mask_bits(p, m0 + r_in_tile, n0, mb);
      if (shared) named_bar_sync(2, HELP_THREADS + EPI_THREADS); else named_bar_sync(1, EPI_THREADS);

      mbar_wait(&tmem_full_bar[buf], (j >> 1) & 1);
      tcgen05_fence_after();

      epilogue_unit(p, &tmOut, taddr, umma_n, tile, split, m0, n0, r_in_tile, lane, wq, s_bias, stage_buf, stores_issued, step,
                    [&]() { if (lane == 0) mbar_arrive(&tmem_empty_bar[buf]); }, shared ? 1 : 0, pre ? &kb : nullptr,
                    pre_mask ? &mb : nullptr);
    }
    if (p.tma_store && lane == 0) tma_store_wait_all();  // smem must outlive the bulk reads
  }

  // teardown: all TMEM reads done before the allocating warp frees the columns
  tcgen05_fence_before();
  __syncthreads();
  if (stamps && threadIdx.x == 0) { stamps[5] = clock64(); stamps[6] = static_cast<long long>(global_timer_ns()); }
  if (warp == 1) tmem_dealloc(tmem_base, TMEM_COLS);
}

// ------------------------------------------------------------------------------------------------
// CTA-pair variant of the 3xTF32 kernel (tcgen05 cta_group::2): a cluster of two CTAs on one TPC owns
// a 256 x bn tile. Each CTA stages its own 128 rows of A and HALF of B, splits them locally, and the
// leader's issuer thread runs M = 256 MMAs that read both CTAs' shared memory. Per SM and K block that
// is 96 KB of shared-memory traffic instead of 144 KB, which lifts the 1-CTA kernel's smem-bandwidth
// bound; L2 -> SM traffic for B halves as well.
//   full[s]   per CTA   : own TMA bytes landed               (producer -> own splitters)
//   ready[s]  leader    : 8 arrivals = 4 splitter warps x 2   (both CTAs' splitters -> leader issuer)
//   empty[s]  per CTA   : tcgen05.commit multicast to both    (issuer -> both producers)
//   tmem_full per CTA   : tcgen05.commit multicast to both    (issuer -> both epilogues)
//   tmem_empty leader   : 8 arrivals                          (both epilogues -> issuer)
template <bool A_MN, bool B_MN, int BK>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NUM_THREADS, 1)
gemm_tf32x3_2cta_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                        const __grid_constant__ CUtensorMap tmOut, const __grid_constant__ CUtensorMap tmA3,
                        const __grid_constant__ CUtensorMap tmB3, const GemmParams p) {
  extern __shared__ uint8_t smem_dyn[];
  pdl_launch_dependents();
  __shared__ __align__(8) uint64_t full_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t ready_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t empty_bar[MAX_STAGES];
  __shared__ __align__(8) uint64_t tmem_full_bar[2];
  __shared__ __align__(8) uint64_t tmem_empty_bar[2];
  __shared__ uint32_t tmem_base_slot;
  __shared__ __align__(16) float s_bias[MAX_BN];

  // warp index / cluster rank through a shuffle: tells the compiler they are warp-uniform, so role code stays in uniform
  // registers (what the MMA issue path needs)
  const int warp = __shfl_sync(0xffffffffu, static_cast<int>(threadIdx.x >> 5), 0);
  const int lane = threadIdx.x & 31;
  const uint32_t rank = __shfl_sync(0xffffffffu, cluster_ctarank(), 0);
  const int pair = blockIdx.x >> 1;
  const int num_pairs = gridDim.x >> 1;

  const uint32_t dyn_base = smem_u32(smem_dyn);
  const uint32_t ring = (dyn_base + 1023u) & ~1023u;
  uint8_t* ring_ptr = smem_dyn + (ring - dyn_base);
  uint8_t* staging = ring_ptr + static_cast<size_t>(p.stages) * p.stage_bytes;

  if (threadIdx.x == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&ready_bar[s], 2 * (SPLIT_THREADS / 32));
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&tmem_full_bar[b], 1);
      mbar_init(&tmem_empty_bar[b], 2 * (EPI_THREADS / 32));
    }
    mbar_fence_init();
  }
  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    if (p.tma_store) tma_prefetch_desc(&tmOut);
    if (p.a3_limit) tma_prefetch_desc(&tmA3);
    if (p.b3_limit) tma_prefetch_desc(&tmB3);
  }
  if (warp == 1) {
    tmem_alloc_2cta(&tmem_base_slot, TMEM_COLS);
    tmem_relinquish_2cta();
  }
  tcgen05_fence_before();
  __syncthreads();
  cluster_sync_all();   // the peer's barriers are initialised before anybody arrives on them remotely
  tcgen05_fence_after();
  pdl_wait();           // everything above ran beside the previous kernel's tail; its results are needed from here on
  const uint32_t tmem_base = __shfl_sync(0xffffffffu, tmem_base_slot, 0);

  const int total_units = p.total_units;
  const int half_bn = p.bn >> 1;
  // RTDC_GEMM_DEBUG_BUF: per-CTA stamps [grid][16]: 0 start clk, 1/2 first-unit epilogue begin/end, 3/4 last-unit epilogue
  // begin/end, 5 end clk, 6/7 end/start ns, 8/9 issuer: first/last unit's MMAs issued
  const bool share_final = !(p.debug & 32) && static_cast<uint32_t>(p.stages) * p.stage_bytes >= 32768u;
  long long* stamps = p.dbg_times ? p.dbg_times + static_cast<size_t>(blockIdx.x) * 16 : nullptr;
  if (stamps && threadIdx.x == 0) { stamps[0] = clock64(); stamps[7] = static_cast<long long>(global_timer_ns()); }

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer (both CTAs)
    // the whole warp walks the loop, one elected lane issues (operands stay in uniform registers, as for the MMAs)
    {
      int s = 0;
      uint32_t ph = 0;
      for (int u = pair; u < total_units; u += num_pairs) {
        const Unit w = decode_unit(p, u);
        const int tile2 = w.tile, kb_begin = w.kb_begin, n_iters = w.n_iters;
        const int m0 = (tile2 % p.m_tiles2) * (2 * BM) + static_cast<int>(rank) * BM;
        const int n0 = (tile2 / p.m_tiles2) * p.bn;
        const int gran2 = B_MN ? 64 : 32;
        const int umma_n = min(p.bn, (min(p.bn, p.N - n0) + gran2 - 1) / gran2 * gran2);   // this tile's MMA width
        const int nb0 = n0 + static_cast<int>(rank) * (umma_n >> 1);                          // this CTA's half of it
        for (int it = 0; it < n_iters; ++it) {
          mbar_wait(&empty_bar[s], ph ^ 1u);
          __syncwarp();
          if (elect_one()) {
            uint8_t* sa = ring_ptr + static_cast<size_t>(s) * p.stage_bytes;
            uint8_t* sb = sa + p.a_bytes;
            if (p.debug & 1) {
              mbar_arrive(&full_bar[s]);
            } else {
              mbar_expect_tx(&full_bar[s], p.a_bytes + p.b_bytes);
              const int k0 = (kb_begin + it) * BK;
              if (!A_MN) {
                tma_load_2d(sa, &tmA, &full_bar[s], k0, m0);
              } else if (m0 + BM <= p.a3_limit) {
                tma_load_3d(sa, &tmA3, &full_bar[s], 0, k0, m0 >> 5);
              } else {
#pragma unroll
                for (int b = 0; b < BM / 32; ++b) tma_load_2d(sa + b * (128 * BK), &tmA, &full_bar[s], m0 + 32 * b, k0);
              }
              if (!B_MN) {
                tma_load_2d(sb, &tmB, &full_bar[s], k0, nb0);
              } else if (nb0 + half_bn <= p.b3_limit) {
                tma_load_3d(sb, &tmB3, &full_bar[s], 0, k0, nb0 >> 5);
              } else {
                for (int b = 0; b < half_bn / 32; ++b) tma_load_2d(sb + b * (128 * BK), &tmB, &full_bar[s], nb0 + 32 * b, k0);
              }
            }
          }
          __syncwarp();
          if (++s == p.stages) { s = 0; ph ^= 1u; }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer (leader CTA only)
    if (rank == 0) {   // the whole warp walks the loop (uniform control flow); one elected lane issues
      const bool a_exact = p.a_exact != 0, b_exact = p.b_exact != 0;
      constexpr uint32_t kmajor_sbo = 8u * BK * 4u;
      constexpr uint32_t mn_box = 128u * BK;
      const uint32_t a_lbo = A_MN ? mn_box : 16u, b_lbo = B_MN ? mn_box : 16u;
      const uint32_t a_sbo = A_MN ? 512u : kmajor_sbo, b_sbo = B_MN ? 512u : kmajor_sbo;
      constexpr uint32_t kmajor_lt = BK == 32 ? UMMA_LAYOUT_SW128 : UMMA_LAYOUT_SW64;
      const uint32_t a_lt = A_MN ? UMMA_LAYOUT_SW128_BASE32B : kmajor_lt;
      const uint32_t b_lt = B_MN ? UMMA_LAYOUT_SW128_BASE32B : kmajor_lt;
      const uint32_t a_kstep = A_MN ? 1024u : 32u, b_kstep = B_MN ? 1024u : 32u;
      int s = 0;
      uint32_t ph = 0;
      int j = 0;
      for (int u = pair; u < total_units; u += num_pairs, ++j) {
        const Unit w = decode_unit(p, u);
        const int tile2 = w.tile, kb_begin = w.kb_begin, n_iters = w.n_iters;
        const int n0 = (tile2 / p.m_tiles2) * p.bn;
        const int gran2 = B_MN ? 64 : 32;
        const int umma_n = min(p.bn, (min(p.bn, p.N - n0) + gran2 - 1) / gran2 * gran2);
        const uint32_t idesc = umma_idesc(UMMA_TF32, A_MN, B_MN, static_cast<uint32_t>(umma_n), 256);
        const int buf = j & 1;
        mbar_wait(&tmem_empty_bar[buf], ((j >> 1) & 1) ^ 1u);
        tcgen05_fence_after();
        const uint32_t tmem_acc = tmem_base + static_cast<uint32_t>(buf) * MAX_BN;
        long long t_wait = 0, t_issue = 0, t_commit = 0;   // issuer time split (stamps only)
        for (int it = 0; it < n_iters; ++it) {
          const long long c0 = stamps ? clock64() : 0;
          mbar_wait(&ready_bar[s], ph);
          __syncwarp();
          if (p.debug & 128) tcgen05_fence_after();   // not needed: the operands are shared memory (TMA tx + proxy-fenced tails), not TMEM
          const long long c1 = stamps ? clock64() : 0;
          const uint32_t sa = ring + static_cast<uint32_t>(s) * p.stage_bytes;
          const uint32_t sb = sa + p.a_bytes;
          const int k_left = p.K - (kb_begin + it) * BK;
          const int ksteps = min(BK / UMMA_K, (k_left + UMMA_K - 1) / UMMA_K);
          if (elect_one()) {
          for (int k = 0; k < ((p.debug & 2) ? 0 : ksteps); ++k) {
            const uint64_t da = umma_smem_desc(sa + k * a_kstep, a_lbo, a_sbo, a_lt);
            const uint64_t db = umma_smem_desc(sb + k * b_kstep, b_lbo, b_sbo, b_lt);
            uint32_t acc = (it > 0 || k > 0) ? 1u : 0u;
            if (!a_exact) {
              const uint64_t da_t = umma_smem_desc(sa + p.a_tail_off + k * a_kstep, a_lbo, a_sbo, a_lt);
              umma_tf32_ss_2cta(tmem_acc, da_t, db, idesc, acc);
              acc = 1u;
            }
            if (!b_exact) {
              const uint64_t db_t = umma_smem_desc(sa + p.b_tail_off + k * b_kstep, b_lbo, b_sbo, b_lt);
              umma_tf32_ss_2cta(tmem_acc, da, db_t, idesc, acc);
              acc = 1u;
            }
            umma_tf32_ss_2cta(tmem_acc, da, db, idesc, acc);
          }
          const long long c2 = stamps ? clock64() : 0;
          tcgen05_commit_2cta(&empty_bar[s], 0x3);        // both CTAs' slots are free once these MMAs have read them
          if (stamps) { const long long c3 = clock64(); t_wait += c1 - c0; t_issue += c2 - c1; t_commit += c3 - c2; }
          }
          __syncwarp();
          if (++s == p.stages) { s = 0; ph ^= 1u; }
        }
        if (elect_one()) {
          tcgen05_commit_2cta(&tmem_full_bar[buf], 0x3);    // both halves of the accumulator are complete
          if (stamps) { if (j == 0) stamps[8] = clock64(); stamps[9] = clock64(); stamps[10] = t_wait; stamps[11] = t_issue; stamps[12] = t_commit; stamps[13] = n_iters; }
        }
        __syncwarp();
      }
    }
    __syncwarp();
  } else if (warp < 2 + SPLIT_THREADS / 32) {
    // ------------------------------------------------------------------ splitters (both CTAs, own tiles)
    const int st = threadIdx.x - 64;
    const uint32_t raw_bytes = p.a_bytes + p.b_bytes;
    const int lo = static_cast<int>(p.split_lo_vec), nvec = static_cast<int>(p.split_hi_vec);
    int s = 0;
    uint32_t ph = 0;
    for (int u = pair; u < total_units; u += num_pairs) {
      const int n_iters = decode_unit(p, u).n_iters;
      for (int it = 0; it < n_iters; ++it) {
        mbar_wait(&full_bar[s], ph);
        const float4* raw = reinterpret_cast<const float4*>(ring_ptr + static_cast<size_t>(s) * p.stage_bytes);
        float4* tail = reinterpret_cast<float4*>(ring_ptr + static_cast<size_t>(s) * p.stage_bytes + raw_bytes) - lo;
        constexpr int SPLIT_ILP = 8;
        for (int base = lo + st; base < ((p.debug & 8) ? 0 : nvec); base += SPLIT_THREADS * SPLIT_ILP) {
          float4 v[SPLIT_ILP];
#pragma unroll
          for (int q = 0; q < SPLIT_ILP; ++q) {
            const int i = base + q * SPLIT_THREADS;
            if (i < nvec) v[q] = raw[i];
          }
#pragma unroll
          for (int q = 0; q < SPLIT_ILP; ++q) {
            const int i = base + q * SPLIT_THREADS;
            if (i < nvec) {
              float4 t;
              t.x = v[q].x - __uint_as_float(__float_as_uint(v[q].x) & 0xFFFFE000u);
              t.y = v[q].y - __uint_as_float(__float_as_uint(v[q].y) & 0xFFFFE000u);
              t.z = v[q].z - __uint_as_float(__float_as_uint(v[q].z) & 0xFFFFE000u);
              t.w = v[q].w - __uint_as_float(__float_as_uint(v[q].w) & 0xFFFFE000u);
              tail[i] = t;
            }
          }
        }
        fence_proxy_async_smem();
        __syncwarp();
        if (lane == 0) mbar_arrive_cluster(&ready_bar[s], 0);   // the leader's issuer waits for all 8 splitter warps
        if (++s == p.stages) { s = 0; ph ^= 1u; }
      }
    }
    // nothing left to split: take the second half of this CTA's final epilogue (staging boxes in the idle ring)
    if (share_final && warp < 2 + HELP_THREADS / 32) {
      const int n_mine = (total_units - pair + num_pairs - 1) / num_pairs;
      const int j = n_mine - 1, u = pair + j * num_pairs;
      const Unit w = decode_unit(p, u);
      const int tile2 = w.tile, split = w.split;
      const int m_tile128 = (tile2 % p.m_tiles2) * 2 + static_cast<int>(rank);
      const int n_tile = tile2 / p.m_tiles2;
      const int m0 = m_tile128 * BM, n0 = n_tile * p.bn;
      const int gran2 = B_MN ? 64 : 32;
      const int umma_n = min(p.bn, (min(p.bn, p.N - n0) + gran2 - 1) / gran2 * gran2);
      const int tile128 = m_tile128 < p.m_tiles ? n_tile * p.m_tiles + m_tile128 : -1;
      const int buf = j & 1;
      const int wq = warp & 3;
      const uint32_t taddr = tmem_base + static_cast<uint32_t>(buf) * MAX_BN + (static_cast<uint32_t>(wq * 32) << 16);
      uint32_t step = p.step_imm;
      if (p.drop_thr16 != 0u && p.step_ptr != nullptr) step = *p.step_ptr;
      uint32_t stores_issued = 0;
      KeepBits hmb;
      if (p.mask_bits_in != nullptr) load_mask_words(p, m0 + wq * 32 + lane, n0, hmb);
      named_bar_sync(2, HELP_THREADS + EPI_THREADS);   // the epilogue warps have staged this unit's bias slice
      mbar_wait(&tmem_full_bar[buf], (j >> 1) & 1);     // the pair's MMAs are done: accumulators complete, both rings idle
      tcgen05_fence_after();
      epilogue_unit(p, &tmOut, taddr, umma_n, tile128, split, m0, n0, wq * 32 + lane, lane, wq, s_bias,
                    ring_ptr + static_cast<size_t>(wq) * 8192, stores_issued, step, []() {}, 2, nullptr,
                    p.mask_bits_in != nullptr ? &hmb : nullptr);
      if (p.tma_store && lane == 0) tma_store_wait_all();
    }
  } else {
    // ------------------------------------------------------------------ epilogue warps (both CTAs, own 128 rows)
    const int et = threadIdx.x - (64 + SPLIT_THREADS);
    const int wq = warp & 3;
    const int r_in_tile = wq * 32 + lane;
    uint32_t step = p.step_imm;
    if (p.drop_thr16 != 0u && p.step_ptr != nullptr) step = *p.step_ptr;
    uint8_t* stage_buf = staging + static_cast<size_t>(wq) * 8192;
    uint32_t stores_issued = 0;
    int j = 0;
    for (int u = pair; u < total_units; u += num_pairs, ++j) {
      const Unit w = decode_unit(p, u);
      const int tile2 = w.tile, split = w.split;
      const int m_tile128 = (tile2 % p.m_tiles2) * 2 + static_cast<int>(rank);
      const int n_tile = tile2 / p.m_tiles2;
      const int m0 = m_tile128 * BM, n0 = n_tile * p.bn;
      const int gran2 = B_MN ? 64 : 32;
      const int umma_n = min(p.bn, (min(p.bn, p.N - n0) + gran2 - 1) / gran2 * gran2);
      const int tile128 = m_tile128 < p.m_tiles ? n_tile * p.m_tiles + m_tile128 : -1;   // -1: rows entirely past M
      const int buf = j & 1;
      const uint32_t taddr = tmem_base + static_cast<uint32_t>(buf) * MAX_BN + (static_cast<uint32_t>(wq * 32) << 16);
      const bool shared = share_final && u + num_pairs >= total_units;   // final unit: split with the splitter warps
      named_bar_sync(1, EPI_THREADS);
      for (int c = et; c < MAX_BN; c += EPI_THREADS)
        s_bias[c] = (p.bias != nullptr && c < p.bn && n0 + c < p.N) ? __ldg(p.bias + n0 + c) : 0.f;
      // dropout flags first (idle time), then the barrier: on the final unit it also waits for the splitter warps
      KeepBits kb;
      const bool pre = p.drop_thr16 != 0u && p.splits == 1;
      if (pre) precompute_keep_bits(p, m0 + r_in_tile, n0, step, kb);
      KeepBits mb;
      const bool pre_mask = (p.mask_src != nullptr || p.mask_bits_in != nullptr) && p.splits == 1;
      if (p.mask_bits_in != nullptr) load_mask_words(p, m0 + r_in_tile, n0, mb);
      else if (pre_mask) precompute_mask_bits(p, m0 + r_in_tile, n0, mb);
      if (shared) named_bar_sync(2, HELP_THREADS + EPI_THREADS); else named_bar_sync(1, EPI_THREADS);
      mbar_wait(&tmem_full_bar[buf], (j >> 1) & 1);
      tcgen05_fence_after();
      if (stamps && et == 0) { if (j == 0) stamps[1] = clock64(); stamps[3] = clock64(); }
      epilogue_unit(p, &tmOut, taddr, umma_n, tile128, split, m0, n0, r_in_tile, lane, wq, s_bias, stage_buf, stores_issued, step,
                    [&]() { if (lane == 0) mbar_arrive_cluster(&tmem_empty_bar[buf], 0); }, shared ? 1 : 0, pre ? &kb : nullptr,
                    pre_mask ? &mb : nullptr);
      if (stamps && et == 0) { if (j == 0) stamps[2] = clock64(); stamps[4] = clock64(); }
    }
    if (p.tma_store && lane == 0) tma_store_wait_all();
  }

  // teardown: nobody may leave (or free TMEM) while the peer can still read this CTA's smem / arrive here
  tcgen05_fence_before();
  __syncthreads();
  if (stamps && threadIdx.x == 0) { stamps[5] = clock64(); stamps[6] = static_cast<long long>(global_timer_ns()); }
  cluster_sync_all();
  if (warp == 1) tmem_dealloc_2cta(tmem_base, TMEM_COLS);
}

// Sums the split-K partial tiles (in split order, so the result is reproducible) and writes the
// output with bias / ReLU applied. One block = 32 rows x 32 columns of one tile chunk; thread =
// one float4, so both the scratch reads and the output writes are full 128-byte lines.
__global__ void __launch_bounds__(256)
splitk_reduce_kernel(const float* __restrict__ scratch, int splits_full, int splits_last, int n_tiles, int m_tiles, int bn, int M, int N,
                     int n_store, float* __restrict__ out, int ldo, const float* __restrict__ bias, int relu, float alpha,
                     const float* __restrict__ row_bias, float row_bias_scale) {
  pdl_launch_dependents();
  pdl_wait();
  const int chunks = (bn + 31) / 32;
  const int tile = blockIdx.x / chunks, c = blockIdx.x % chunks;
  const int m_tile = tile % m_tiles, n_tile = tile / m_tiles;
  const int splits = n_tile == n_tiles - 1 ? splits_last : splits_full;   // partial slots are laid out with stride splits_full
  const int r = blockIdx.y * 32 + (threadIdx.x >> 3), q = threadIdx.x & 7;
  const int row = m_tile * BM + r;
  const int col = n_tile * bn + c * 32 + q * 4;
  if (row >= M || col >= n_store || c * 32 + q * 4 >= bn) return;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  const float* src = scratch + (static_cast<size_t>(tile) * splits_full) * (BM * MAX_BN) + (static_cast<size_t>(c) * BM + r) * 32 + q * 4;
  if (col < N) {
    constexpr int ZI = 8;   // loads in flight; the sum is still formed in split order
    for (int z0 = 0; z0 < splits; z0 += ZI) {
      float4 t[ZI];
#pragma unroll
      for (int u = 0; u < ZI; ++u)
        if (z0 + u < splits) t[u] = __ldcg(reinterpret_cast<const float4*>(src + static_cast<size_t>(z0 + u) * (BM * MAX_BN)));
#pragma unroll
      for (int u = 0; u < ZI; ++u)
        if (z0 + u < splits) { acc.x += t[u].x; acc.y += t[u].y; acc.z += t[u].z; acc.w += t[u].w; }
    }
  }
  float v[4] = {acc.x, acc.y, acc.z, acc.w};
  const float rb = row_bias != nullptr ? __ldg(row_bias + row) * row_bias_scale : 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float x = 0.f;
    if (col + j < N) {
      x = fmaf(v[j], alpha, bias != nullptr ? __ldg(bias + col + j) : 0.f) + rb;
      if (relu) x = fmaxf(x, 0.f);
    }
    v[j] = x;
  }
  float* dst = out + static_cast<size_t>(row) * ldo + col;
  if (col + 3 < n_store && ((reinterpret_cast<uintptr_t>(dst) & 15u) == 0)) {
    *reinterpret_cast<float4*>(dst) = make_float4(v[0], v[1], v[2], v[3]);
  } else {
#pragma unroll
    for (int j = 0; j < 4; ++j)
      if (col + j < n_store) dst[j] = v[j];
  }
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
                    CUtensorMapSwizzle swizzle) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return fail(RTDC_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  if ((reinterpret_cast<uintptr_t>(base) & 15u) != 0) return fail(RTDC_ERR_ARG, "operand pointer must be 16-byte aligned");
  if ((static_cast<size_t>(ld) * 4) % 16 != 0) return fail(RTDC_ERR_ARG, "operand leading dimension must be a multiple of 4 floats (got %d)", ld);
  cuuint64_t dims[2] = {static_cast<cuuint64_t>(inner), static_cast<cuuint64_t>(outer)};
  cuuint64_t strides[1] = {static_cast<cuuint64_t>(ld) * 4};
  cuuint32_t box[2] = {static_cast<cuuint32_t>(box_inner), static_cast<cuuint32_t>(box_outer)};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, swizzle,
                   CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(RTDC_ERR_CUDA, "cuTensorMapEncodeTiled failed with CUresult %d", static_cast<int>(r));
  return RTDC_OK;
}

// 3-D view of a row-major [k_extent, ld] fp32 matrix whose first `blocks * 32` columns are whole 32-element blocks:
// (32 contiguous floats, k rows of pitch ld, blocks of pitch 128 B). One box = [box_blocks][box_k][32], which is exactly the
// shared-memory layout of box_blocks consecutive 2-D MN-major boxes.
static int make_map3_mn(CUtensorMap* map, const float* base, int blocks, int k_extent, int ld, int box_k, int box_blocks) {
  EncodeTiledFn enc = get_encode_fn();
  if (!enc) return fail(RTDC_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
  cuuint64_t dims[3] = {32, static_cast<cuuint64_t>(k_extent), static_cast<cuuint64_t>(blocks)};
  cuuint64_t strides[2] = {static_cast<cuuint64_t>(ld) * 4, 128};
  cuuint32_t box[3] = {32, static_cast<cuuint32_t>(box_k), static_cast<cuuint32_t>(box_blocks)};
  cuuint32_t estr[3] = {1, 1, 1};
  CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 3, const_cast<float*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(RTDC_ERR_CUDA, "cuTensorMapEncodeTiled (3-D MN-major view) failed with CUresult %d", static_cast<int>(r));
  return RTDC_OK;
}

struct Plan {
  int two_cta, m_tiles2;
  int a_exact, b_exact;
  int bk;
  int bn, stages, splits, kb_per_split, num_kb, m_tiles, n_tiles;
  int splits_last, kb_per_split_last;   // split count of the last (narrower) N column's tiles, see GemmParams
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

// Tuning knobs from the environment, read ONCE when the library first plans a GEMM (not per launch):
//   RTDC_GEMM_2CTA=0/1        force the 1-CTA / CTA-pair kernel      RTDC_GEMM_2CTA_BK=16   16-wide K blocks for the pair kernel
//   RTDC_GEMM_SPLITS_LAST=n   K splits of a narrow last N column     RTDC_GEMM_NO_TMA3D     2-D maps for MN-major operands
//   RTDC_GEMM_NO_TMA_STORE    direct-store epilogue
// Builds with -DRTDC_GEMM_DEBUG_HOOKS additionally honour RTDC_GEMM_DEBUG (role mask) and RTDC_GEMM_DEBUG_BUF (device pointer of
// a clock-stamp buffer, used by tools/gemm_stamps*.py); the default build ignores both.
struct GemmKnobs {
  int force_2cta = -1;     // -1 unset, 0, 1
  int bk16 = 0;
  int splits_last = 0;     // 0 unset
  int no_tma3d = 0, no_tma_store = 0;
  int debug = 0;
  long long* dbg_times = nullptr;
};
static GemmKnobs read_knobs() {
    GemmKnobs r;
    if (const char* e = getenv("RTDC_GEMM_2CTA")) r.force_2cta = e[0] == '1' ? 1 : (e[0] == '0' ? 0 : -1);
    if (const char* e = getenv("RTDC_GEMM_2CTA_BK")) r.bk16 = atoi(e) == 16;
    if (const char* e = getenv("RTDC_GEMM_SPLITS_LAST")) r.splits_last = atoi(e);
    r.no_tma3d = getenv("RTDC_GEMM_NO_TMA3D") != nullptr;
    r.no_tma_store = getenv("RTDC_GEMM_NO_TMA_STORE") != nullptr;
#ifdef RTDC_GEMM_DEBUG_HOOKS
    if (const char* e = getenv("RTDC_GEMM_DEBUG")) r.debug = atoi(e);
    if (const char* e = getenv("RTDC_GEMM_DEBUG_BUF")) r.dbg_times = reinterpret_cast<long long*>(strtoull(e, nullptr, 0));
#endif
    return r;
}
#ifdef RTDC_GEMM_DEBUG_HOOKS
static GemmKnobs knobs() { return read_knobs(); }                 // probe tools flip the variables between launches
#else
static const GemmKnobs& knobs() {
  static const GemmKnobs k = read_knobs();
  return k;
}
#endif

static int make_plan(const rtdc_gemm_desc* d, Plan* pl, bool need_device) {
  if (!d) return fail(RTDC_ERR_ARG, "null descriptor");
  if (d->M <= 0 || d->N <= 0 || d->K <= 0) return fail(RTDC_ERR_ARG, "M, N, K must be positive (got %d, %d, %d)", d->M, d->N, d->K);
  if (d->precision != 1 && d->precision != 3) return fail(RTDC_ERR_ARG, "precision must be 1 (tf32) or 3 (3xTF32)");
  // CTA pairs (cta_group::2) for the fp32-accurate path: a pair shares one 256-row tile, each CTA stages half of B, so
  // shared-memory traffic per SM drops by a third and the main loop becomes tensor-bound. Used when the tile grid pairs up
  // and the coarser N granularity does not cost more padding than it saves. RTDC_GEMM_2CTA=0/1 forces the choice.
  const GemmKnobs kn = knobs();
  const int gran2 = d->b_mn ? 64 : 32, gran1 = d->b_mn ? 32 : 16;
  bool two_cta = d->precision == 3 && d->M > BM && d->N >= 128;
  if (two_cta && d->bn != 0 && d->bn % gran2 != 0) two_cta = false;
  if (two_cta && d->bn == 0 && kn.force_2cta != 1) {
    const int nt = ceil_div(d->N, MAX_BN);
    const int bn1 = round_up(ceil_div(d->N, nt), (nt > 1 && gran1 < 32) ? 32 : gran1), bn2 = round_up(ceil_div(d->N, nt), gran2);
    if (ceil_div(d->N, bn2) * bn2 * 100 > ceil_div(d->N, bn1) * bn1 * 120) two_cta = false;
  }
  if (kn.force_2cta == 0) two_cta = false;
  pl->two_cta = two_cta ? 1 : 0;
  const int gran = two_cta ? (d->b_mn ? 64 : 32) : (d->b_mn ? 32 : 16);
  int bn = d->bn;
  if (bn == 0) {
    const int nt = ceil_div(d->N, MAX_BN);
    // several N tiles: whole 32-column store boxes per tile, so a tile's last box cannot spill into its neighbour
    bn = round_up(ceil_div(d->N, nt), (nt > 1 && gran < 32) ? 32 : gran);
  }
  if (bn < gran || bn > MAX_BN || bn % gran != 0)
    return fail(RTDC_ERR_ARG, "bn must be a multiple of %d in [%d, %d] (got %d)", gran, gran, MAX_BN, bn);
  pl->bn = bn;
  pl->m_tiles = ceil_div(d->M, BM);
  pl->m_tiles2 = ceil_div(d->M, 2 * BM);
  pl->n_tiles = ceil_div(d->N, bn);
  if (d->n_store > d->N && d->n_store > pl->n_tiles * bn)
    return fail(RTDC_ERR_ARG, "n_store (%d) reaches beyond the last N tile", d->n_store);
  // K block: the 1-CTA 3xTF32 kernel needs 16 to fit four stages of raw + tail tiles; a CTA pair stages half of B, so 32
  // fits three (and halves the per-stage barrier / commit overhead of the issuer thread). RTDC_GEMM_2CTA_BK=16 forces 16.
  pl->bk = d->precision == 3 ? ((two_cta && !kn.bk16) ? 32 : 16) : 32;
  pl->num_kb = ceil_div(d->K, pl->bk);
  pl->a_bytes = BM * pl->bk * 4;
  pl->b_bytes = static_cast<uint32_t>(two_cta ? bn / 2 : bn) * pl->bk * 4;   // a CTA of a pair stages half of B
  // an operand declared exact needs no tail tile (3xTF32 kernels; the 1-pass kernel has no tails at all)
  pl->a_exact = (d->precision == 3 && d->a_exact) ? 1 : 0;
  pl->b_exact = (d->precision == 3 && d->b_exact) ? 1 : 0;
  pl->stage_bytes = pl->a_bytes + pl->b_bytes;
  if (d->precision == 3) pl->stage_bytes += (pl->a_exact ? 0u : pl->a_bytes) + (pl->b_exact ? 0u : pl->b_bytes);
  int stages = static_cast<int>((SMEM_LIMIT - STATIC_SMEM_RESERVE - 1024 - EPI_STAGING_BYTES) / pl->stage_bytes);
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages > pl->num_kb) stages = pl->num_kb < 1 ? 1 : pl->num_kb;
  if (stages < 1) return fail(RTDC_ERR_LIMIT, "tile does not fit shared memory");
  pl->stages = stages;
  pl->smem = static_cast<size_t>(stages) * pl->stage_bytes + EPI_STAGING_BYTES + 1024;
  const int tiles = (two_cta ? pl->m_tiles2 : pl->m_tiles) * pl->n_tiles;   // schedulable tiles (pair tiles in 2-CTA mode)
  int splits = d->splits;
  if (splits == 0) {
    splits = 1;
    const int sms = (need_device ? num_sms() : 148) / (two_cta ? 2 : 1);
    if (tiles * 2 <= sms && d->K >= 512 && d->drop_p <= 0.f && d->mask_src == nullptr && d->mask_bits_in == nullptr && d->mask_bits_out == nullptr) {
      splits = sms / tiles;
      const int max_by_k = d->K / 256;  // keep >= 256 K elements per split
      if (splits > max_by_k) splits = max_by_k;
      if (splits > 64) splits = 64;
      if (splits < 1) splits = 1;
    }
  }
  if (splits > 1 && (d->drop_p > 0.f || d->mask_src != nullptr || d->colsum_partial != nullptr || d->mask_bits_in != nullptr ||
                     d->mask_bits_out != nullptr)) splits = 1;  // those epilogues need the full sum in TMEM
  if (splits > pl->num_kb) splits = pl->num_kb;
  pl->kb_per_split = ceil_div(pl->num_kb, splits);
  pl->splits = ceil_div(pl->num_kb, pl->kb_per_split);  // no empty split
  pl->splits_last = pl->splits; pl->kb_per_split_last = pl->kb_per_split;
  // Weighted split-K: when the last N column is narrower than the others its units are cheaper, so it gets fewer splits and
  // the freed CTAs shorten everybody else's K range. Cost of a unit ~ K blocks x (fixed per-stage work + MMA width).
  if (d->splits == 0 && pl->splits > 1 && pl->n_tiles > 1) {
    const int w_last = round_up(d->N - (pl->n_tiles - 1) * bn, gran) < bn ? round_up(d->N - (pl->n_tiles - 1) * bn, gran) : bn;
    const int m_sched = two_cta ? pl->m_tiles2 : pl->m_tiles;
    const int sms = (need_device ? num_sms() : 148) / (two_cta ? 2 : 1);
    const int slots = sms / m_sched, nt = pl->n_tiles;
    int cap = d->K / 256 < 64 ? d->K / 256 : 64;
    if (cap < 1) cap = 1;
    if (w_last < bn && slots >= nt) {
      const double c_full = 1.0, c_last = 0.75 + 0.25 * static_cast<double>(w_last) / bn;   // measured: a 64-wide unit costs ~80 % of a 256-wide one per K block
      double best = 1e30;
      int best_f = pl->splits, best_l = pl->splits;
      for (int sl = 1; sl <= cap && sl <= slots - (nt - 1); ++sl) {
        int sf = (slots - sl) / (nt - 1);
        if (sf > cap) sf = cap;
        if (sf < sl) continue;                       // partial slots are laid out with stride splits_full
        if (kn.splits_last && kn.splits_last != sl) continue;
        const double cost_f = ceil_div(pl->num_kb, sf) * c_full, cost_l = ceil_div(pl->num_kb, sl) * c_last;
        const double cost = cost_f > cost_l ? cost_f : cost_l;
        if (cost <= best + 1e-9) { best = cost; best_f = sf; best_l = sl; }   // ties: more splits for the last column
      }
      pl->kb_per_split = ceil_div(pl->num_kb, best_f);
      pl->splits = ceil_div(pl->num_kb, pl->kb_per_split);
      pl->kb_per_split_last = ceil_div(pl->num_kb, best_l);
      pl->splits_last = ceil_div(pl->num_kb, pl->kb_per_split_last);
      if (pl->splits_last > pl->splits) { pl->splits_last = pl->splits; pl->kb_per_split_last = pl->kb_per_split; }
    }
  }
  return RTDC_OK;
}

// scratch layout: [tiles x splits] partial tiles of BM x MAX_BN floats, chunk-major inside a tile
static size_t scratch_bytes_for(const Plan& pl) {
  if (pl.splits <= 1) return 0;
  const size_t tiles = static_cast<size_t>(pl.m_tiles) * pl.n_tiles;
  return tiles * pl.splits * (static_cast<size_t>(BM) * MAX_BN * 4);
}

template <bool A_MN, bool B_MN, int BK>
static int launch_2cta_bk(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const CUtensorMap& ta3, const CUtensorMap& tb3,
                       const GemmParams& gp, const Plan& pl, cudaStream_t st) {
  auto kern = gemm_tf32x3_2cta_kernel<A_MN, B_MN, BK>;
  RTDC_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT - STATIC_SMEM_RESERVE));
  const int units = gp.total_units;
  const int pairs = units < num_sms() / 2 ? units : num_sms() / 2;   // persistent: one CTA pair per TPC
  RTDC_CUDA_CHECK(launch_pdl(kern, dim3(2 * pairs), dim3(NUM_THREADS), pl.smem, st, ta, tb, to, ta3, tb3, gp));
  return RTDC_OK;
}

template <bool A_MN, bool B_MN>
static int launch_2cta(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const CUtensorMap& ta3, const CUtensorMap& tb3,
                       const GemmParams& gp, const Plan& pl, cudaStream_t st) {
  return pl.bk == 32 ? launch_2cta_bk<A_MN, B_MN, 32>(ta, tb, to, ta3, tb3, gp, pl, st)
                     : launch_2cta_bk<A_MN, B_MN, 16>(ta, tb, to, ta3, tb3, gp, pl, st);
}

template <bool A_MN, bool B_MN, int NSPLIT, int BK>
static int launch(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to, const CUtensorMap& ta3, const CUtensorMap& tb3,
                  const GemmParams& gp, const Plan& pl, cudaStream_t st) {
  auto kern = gemm_tf32_kernel<A_MN, B_MN, NSPLIT, BK>;
  RTDC_CUDA_CHECK(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM_LIMIT - STATIC_SMEM_RESERVE));
  const int units = gp.total_units;
  const int grid = units < num_sms() ? units : num_sms();   // persistent: one CTA per SM walks the work units
  RTDC_CUDA_CHECK(launch_pdl(kern, dim3(grid), dim3(NUM_THREADS), pl.smem, st, ta, tb, to, ta3, tb3, gp));
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

static int dispatch(const rtdc_gemm_desc* d, bool three, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to,
                    const CUtensorMap& ta3, const CUtensorMap& tb3, const rtdc::GemmParams& gp, const rtdc::Plan& pl, cudaStream_t st);

RTDC_API int rtdc_gemm(const rtdc_gemm_desc* d, void* stream) {
  Plan pl;
  int rc = make_plan(d, &pl, true);
  if (rc != RTDC_OK) return rc;
  if (!d->a || !d->b || !d->out) return fail(RTDC_ERR_ARG, "a, b and out must be non-null");
  if (d->drop_p < 0.f || d->drop_p >= 1.f) return fail(RTDC_ERR_ARG, "drop_p must be in [0, 1)");
  const int n_store = d->n_store > d->N ? d->n_store : d->N;
  if (d->ldo < n_store) return fail(RTDC_ERR_ARG, "ldo (%d) smaller than the stored width (%d)", d->ldo, n_store);

  CUtensorMap ta, tb;
  const CUtensorMapSwizzle sw_k = pl.bk == 32 ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_64B;
  const CUtensorMapSwizzle sw_mn = CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B;
  if (!d->a_mn) rc = make_map(&ta, d->a, d->K, d->M, d->lda, pl.bk, BM, sw_k);
  else          rc = make_map(&ta, d->a, d->M, d->K, d->lda, 32, pl.bk, sw_mn);
  if (rc != RTDC_OK) return rc;
  if (!d->b_mn) rc = make_map(&tb, d->b, d->K, d->N, d->ldb, pl.bk, pl.two_cta ? pl.bn / 2 : pl.bn, sw_k);
  else          rc = make_map(&tb, d->b, d->N, d->K, d->ldb, 32, pl.bk, sw_mn);
  if (rc != RTDC_OK) return rc;

  GemmParams gp;
  memset(&gp, 0, sizeof(gp));
  // one-instruction loads of MN-major tiles over the whole 32-element blocks of the operand (RTDC_GEMM_NO_TMA3D=1: off)
  CUtensorMap ta3, tb3;
  memset(&ta3, 0, sizeof(ta3)); memset(&tb3, 0, sizeof(tb3));
  const bool use3 = !knobs().no_tma3d;
  if (use3 && d->a_mn && d->M >= BM) {
    rc = make_map3_mn(&ta3, d->a, d->M / 32, d->K, d->lda, pl.bk, BM / 32);
    if (rc != RTDC_OK) return rc;
    gp.a3_limit = (d->M / 32) * 32;
  }
  const int b_cols = pl.two_cta ? pl.bn / 2 : pl.bn;   // columns of B one CTA stages
  if (use3 && d->b_mn && d->N >= b_cols) {
    rc = make_map3_mn(&tb3, d->b, d->N / 32, d->K, d->ldb, pl.bk, b_cols / 32);
    if (rc != RTDC_OK) return rc;
    gp.b3_limit = (d->N / 32) * 32;
  }
  gp.M = d->M; gp.N = d->N; gp.K = d->K;
  gp.bn = pl.bn; gp.stages = pl.stages; gp.splits = pl.splits; gp.kb_per_split = pl.kb_per_split; gp.num_kb = pl.num_kb;
  gp.m_tiles = pl.m_tiles; gp.n_tiles = pl.n_tiles; gp.m_tiles2 = pl.m_tiles2;
  gp.m_sched = pl.two_cta ? pl.m_tiles2 : pl.m_tiles;
  gp.splits_last = pl.splits_last; gp.kb_per_split_last = pl.kb_per_split_last;
  if (pl.splits_last != pl.splits) {
    gp.tiles_full = gp.m_sched * (pl.n_tiles - 1);
    gp.full_units = gp.tiles_full * pl.splits;
    gp.total_units = gp.full_units + gp.m_sched * pl.splits_last;
  } else {   // uniform: one numbering over all tiles
    gp.tiles_full = gp.m_sched * pl.n_tiles;
    gp.full_units = gp.tiles_full * pl.splits;
    gp.total_units = gp.full_units;
  }
  gp.a_bytes = pl.a_bytes; gp.b_bytes = pl.b_bytes; gp.stage_bytes = pl.stage_bytes;
  gp.a_exact = pl.a_exact; gp.b_exact = pl.b_exact;
  gp.a_tail_off = pl.a_bytes + pl.b_bytes;
  gp.b_tail_off = pl.a_bytes + pl.b_bytes + (pl.a_exact ? 0u : pl.a_bytes);
  gp.split_lo_vec = pl.a_exact ? pl.a_bytes >> 4 : 0u;
  gp.split_hi_vec = pl.b_exact ? pl.a_bytes >> 4 : (pl.a_bytes + pl.b_bytes) >> 4;
  gp.alpha = d->alpha == 0.f ? 1.f : d->alpha;
  gp.row_bias = d->row_bias; gp.row_bias_scale = d->row_bias_scale;
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
  if (d->mask_bits_in != nullptr && d->mask_src != nullptr) return fail(RTDC_ERR_ARG, "give the backward mask either as mask_src or as mask_bits_in, not both");
  if ((d->mask_bits_in != nullptr && d->ld_mask_in * 32 < d->N) || (d->mask_bits_out != nullptr && d->ld_mask_out * 32 < d->N))
    return fail(RTDC_ERR_ARG, "mask bit rows need at least ceil(N/32) words");
  if ((d->mask_bits_in != nullptr || d->mask_bits_out != nullptr) && pl.n_tiles > 1 && pl.bn % 32 != 0)
    return fail(RTDC_ERR_ARG, "bit-packed masks need N tiles that are multiples of 32 columns");
  gp.mask_bits_out = d->mask_bits_out; gp.ld_mask_out = d->ld_mask_out;
  gp.mask_bits_in = d->mask_bits_in; gp.ld_mask_in = d->ld_mask_in;
  if (pl.splits > 1) {
    const size_t need = scratch_bytes_for(pl);
    if (!d->scratch || d->scratch_bytes < need)
      return fail(RTDC_ERR_ARG, "split-K x%d needs %zu bytes of scratch (got %zu)", pl.splits, need, d->scratch_bytes);
    gp.scratch = d->scratch;
  }
  // output through TMA stores when the destination allows a tensor map (16-byte aligned base and
  // row pitch); otherwise the epilogue falls back to direct per-thread stores
  CUtensorMap to;
  memset(&to, 0, sizeof(to));
  if ((reinterpret_cast<uintptr_t>(d->out) & 15u) == 0 && (static_cast<size_t>(d->ldo) * 4) % 16 == 0 &&
      (pl.bn % 32 == 0 || pl.n_tiles == 1) && !knobs().no_tma_store) {
    rc = make_map(&to, d->out, n_store, d->M, d->ldo, 32, 32, CU_TENSOR_MAP_SWIZZLE_128B);
    if (rc != RTDC_OK) return rc;
    gp.tma_store = 1;
  }
  if (d->colsum_partial != nullptr) {
    if (!gp.tma_store || pl.splits > 1) return fail(RTDC_ERR_ARG, "colsum_partial needs an aligned output (TMA-store epilogue) and no K split");
    if (d->ld_colsum < d->N) return fail(RTDC_ERR_ARG, "ld_colsum (%d) smaller than N (%d)", d->ld_colsum, d->N);
    gp.colsum_partial = d->colsum_partial; gp.ld_colsum = d->ld_colsum;
  }
  gp.debug = knobs().debug;             // both stay 0 / null unless built with -DRTDC_GEMM_DEBUG_HOOKS
  gp.dbg_times = knobs().dbg_times;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const bool three = d->precision == 3;
  rc = dispatch(d, three, ta, tb, to, ta3, tb3, gp, pl, st);
  if (rc != RTDC_OK) return rc;
  if (pl.splits > 1) {
    dim3 grid(pl.m_tiles * pl.n_tiles * ((pl.bn + 31) / 32), BM / 32);
    RTDC_CUDA_CHECK(launch_pdl(splitk_reduce_kernel, grid, dim3(256), 0, st, gp.scratch, pl.splits, pl.splits_last, pl.n_tiles, pl.m_tiles, pl.bn,
                               d->M, d->N, n_store, d->out, d->ldo, d->bias, d->relu, gp.alpha, d->row_bias, d->row_bias_scale));
  }
  return RTDC_OK;
}

static int dispatch(const rtdc_gemm_desc* d, bool three, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& to,
                    const CUtensorMap& ta3, const CUtensorMap& tb3, const rtdc::GemmParams& gp, const rtdc::Plan& pl, cudaStream_t st) {
  if (pl.two_cta) {
    if (!d->a_mn && !d->b_mn) return launch_2cta<false, false>(ta, tb, to, ta3, tb3, gp, pl, st);
    if (!d->a_mn && d->b_mn)  return launch_2cta<false, true>(ta, tb, to, ta3, tb3, gp, pl, st);
    if (d->a_mn && !d->b_mn)  return launch_2cta<true, false>(ta, tb, to, ta3, tb3, gp, pl, st);
    return launch_2cta<true, true>(ta, tb, to, ta3, tb3, gp, pl, st);
  }
  if (!d->a_mn && !d->b_mn) return three ? launch<false, false, 3, 16>(ta, tb, to, ta3, tb3, gp, pl, st) : launch<false, false, 1, 32>(ta, tb, to, ta3, tb3, gp, pl, st);
  if (!d->a_mn && d->b_mn)  return three ? launch<false, true, 3, 16>(ta, tb, to, ta3, tb3, gp, pl, st)  : launch<false, true, 1, 32>(ta, tb, to, ta3, tb3, gp, pl, st);
  if (d->a_mn && !d->b_mn)  return three ? launch<true, false, 3, 16>(ta, tb, to, ta3, tb3, gp, pl, st)  : launch<true, false, 1, 32>(ta, tb, to, ta3, tb3, gp, pl, st);
  return three ? launch<true, true, 3, 16>(ta, tb, to, ta3, tb3, gp, pl, st) : launch<true, true, 1, 32>(ta, tb, to, ta3, tb3, gp, pl, st);
}
