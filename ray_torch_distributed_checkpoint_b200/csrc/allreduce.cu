// Collectives over peer-mapped symmetric memory (csrc/symm_mem.cpp): the DDP gradient-bucket
// all-reduce of SURVEY.md §2.5-C3 with the 1/world scale (and bf16<->fp32 accumulate) fused in,
// plus broadcast (C2/C4) and barrier (C5). No NCCL on this path.
//
// Protocol. Block b of every rank forms a group with block b of every other rank. A group meets at
// a barrier by bumping per-(block, source-rank) arrival counters in the peers' signal pads with
// st.release.sys and spinning (bounded) on its own pad with ld.acquire.sys; counters only grow, the
// expected value comes from a per-block epoch kept in local memory, so nothing is ever reset and a
// captured CUDA graph can replay the kernel. Work is partitioned identically on all ranks, so block
// b only ever needs data that block b of its peers produced.
//
//   one-shot : barrier; every rank loads all peers' elements and writes the sum to a local output.
//              (world-1) x count inbound per GPU, one round -> small messages.
//   two-shot : barrier; rank r sums slice r from all peers (rank order => bit-identical everywhere)
//              and stores the result into slice r of EVERY peer (push all-gather); barrier.
//              2 x (world-1)/world x count per direction.
//   NVLS     : same schedule, but one multimem.ld_reduce pulls the switch-reduced value of slice r
//              and one multimem.st broadcasts the result: ~count/world in and out of each SM.
#include <cuda_bf16.h>

#include "common.h"
#include "ptx.cuh"
#include "symm.h"

namespace rtdc {

constexpr int AR_THREADS = 512;

struct SymmView {  // mirrors the device tables of struct Symm
  void** peers;
  void** pads;
  unsigned int* epochs;
  void* mc;
  int rank, world;
};

__device__ __forceinline__ void st_release_sys(unsigned int* p, unsigned int v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ unsigned int ld_acquire_sys(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// Barrier among block `blockIdx.x` of all ranks. `seq` = value the counters must reach.
__device__ __forceinline__ void group_barrier(const SymmView& sv, unsigned int seq) {
  __syncthreads();  // everything this block wrote is ordered before the release below
  if (threadIdx.x < sv.world) {
    const int peer = threadIdx.x;
    unsigned int* theirs = static_cast<unsigned int*>(sv.pads[peer]) + blockIdx.x * RTDC_SYMM_MAX_WORLD + sv.rank;
    st_release_sys(theirs, seq);
    const unsigned int* mine = static_cast<const unsigned int*>(sv.pads[sv.rank]) + blockIdx.x * RTDC_SYMM_MAX_WORLD + peer;
    if (static_cast<int>(ld_acquire_sys(mine) - seq) < 0) {
      const long long t0 = clock64();
      while (static_cast<int>(ld_acquire_sys(mine) - seq) < 0) {
        if (clock64() - t0 > RTDC_WAIT_TIMEOUT_CYCLES * 4) {
          printf("rtdc: rank %d block %d timed out waiting for rank %d (seq %u)\n", sv.rank, blockIdx.x, peer, seq);
          __trap();
        }
      }
    }
  }
  __syncthreads();
}

// Each launch uses `rounds` barriers; the block's epoch counts launches-worth of rounds so far.
__device__ __forceinline__ unsigned int epoch_begin(const SymmView& sv) { return sv.epochs[blockIdx.x]; }
__device__ __forceinline__ void epoch_end(const SymmView& sv, unsigned int e, unsigned int rounds) {
  __syncthreads();
  if (threadIdx.x == 0) sv.epochs[blockIdx.x] = e + rounds;
}

struct alignas(16) Vec4 { float x, y, z, w; };

template <typename T> struct Pack;  // 16 bytes of T <-> floats
template <> struct Pack<float> {
  static constexpr int N = 4;
  __device__ static void load(const void* p, float (&f)[4]) {
    const float4 v = *reinterpret_cast<const float4*>(p);
    f[0] = v.x; f[1] = v.y; f[2] = v.z; f[3] = v.w;
  }
  __device__ static void store(void* p, const float (&f)[4]) { *reinterpret_cast<float4*>(p) = make_float4(f[0], f[1], f[2], f[3]); }
};
template <> struct Pack<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ static void load(const void* p, float (&f)[8]) {
    const uint4 v = *reinterpret_cast<const uint4*>(p);
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      f[2 * i] = __uint_as_float(w[i] << 16);
      f[2 * i + 1] = __uint_as_float(w[i] & 0xFFFF0000u);
    }
  }
  __device__ static void store(void* p, const float (&f)[8]) {
    uint32_t w[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(f[2 * i], f[2 * i + 1]);
      w[i] = *reinterpret_cast<const uint32_t*>(&h);
    }
    *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
  }
};

// peer loads bypass L1 (another GPU may rewrite the line between launches)
__device__ __forceinline__ uint4 ld_peer16(const void* p) {
  uint4 v;
  asm volatile("ld.global.relaxed.sys.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void st_peer16(void* p, const uint4& v) {
  asm volatile("st.global.relaxed.sys.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <typename T, int WORLD>
__device__ __forceinline__ void reduce_vec(const SymmView& sv, size_t byte_off, float scale, uint4& out) {
  constexpr int N = Pack<T>::N;
  uint4 raw[WORLD];
#pragma unroll
  for (int p = 0; p < WORLD; ++p) raw[p] = ld_peer16(static_cast<const char*>(sv.peers[p]) + byte_off);  // all loads in flight
  float acc[N];
#pragma unroll
  for (int i = 0; i < N; ++i) acc[i] = 0.f;
#pragma unroll
  for (int p = 0; p < WORLD; ++p) {  // fixed rank order
    float f[N];
    Pack<T>::load(&raw[p], f);
#pragma unroll
    for (int i = 0; i < N; ++i) acc[i] += f[i];
  }
#pragma unroll
  for (int i = 0; i < N; ++i) acc[i] *= scale;
  Pack<T>::store(&out, acc);
}

// ------------------------------------------------------------------ two-shot (in place)
constexpr int AR_UNROLL = 4;   // 16-byte vectors per thread per trip (NVLS); the P2P kernels hold UNROLL x WORLD vectors in registers
template <int WORLD> struct P2PUnroll { static constexpr int value = WORLD <= 4 ? 4 : 2; };

template <typename T, int WORLD>
__global__ void __launch_bounds__(AR_THREADS)
two_shot_kernel(SymmView sv, size_t offset, size_t nvec_total, float scale) {
  constexpr int U = P2PUnroll<WORLD>::value;
  const unsigned int e = epoch_begin(sv);
  group_barrier(sv, e + 1);  // every rank's contribution is in its buffer
  // slice owned by this rank, in 16-byte vectors
  const size_t per = (nvec_total + WORLD - 1) / WORLD;
  const size_t begin = min(nvec_total, per * sv.rank), end = min(nvec_total, begin + per);
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t v0 = begin + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < end; v0 += stride * U) {
    uint4 raw[U][WORLD];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + u * stride;
      if (v < end) {
#pragma unroll
        for (int p = 0; p < WORLD; ++p) raw[u][p] = ld_peer16(static_cast<const char*>(sv.peers[p]) + offset + v * 16);
      }
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + u * stride;
      if (v >= end) continue;
      constexpr int N = Pack<T>::N;
      float acc[N];
#pragma unroll
      for (int i = 0; i < N; ++i) acc[i] = 0.f;
#pragma unroll
      for (int p = 0; p < WORLD; ++p) {  // fixed rank order
        float f[N];
        Pack<T>::load(&raw[u][p], f);
#pragma unroll
        for (int i = 0; i < N; ++i) acc[i] += f[i];
      }
#pragma unroll
      for (int i = 0; i < N; ++i) acc[i] *= scale;
      uint4 r;
      Pack<T>::store(&r, acc);
      const size_t off = offset + v * 16;
#pragma unroll
      for (int p = 0; p < WORLD; ++p) st_peer16(static_cast<char*>(sv.peers[(sv.rank + p) % WORLD]) + off, r);
    }
  }
  group_barrier(sv, e + 2);  // all pushes into my buffer have landed
  epoch_end(sv, e, 2);
}

// ------------------------------------------------------------------ one-shot (out of place)
template <typename T, int WORLD>
__global__ void __launch_bounds__(AR_THREADS)
one_shot_kernel(SymmView sv, size_t offset, size_t nvec_total, float scale, void* out_local) {
  const unsigned int e = epoch_begin(sv);
  group_barrier(sv, e + 1);
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec_total;
       v += static_cast<size_t>(gridDim.x) * blockDim.x) {
    uint4 r;
    reduce_vec<T, WORLD>(sv, offset + v * 16, scale, r);
    reinterpret_cast<uint4*>(out_local)[v] = r;
  }
  group_barrier(sv, e + 2);  // nobody rewrites its input while a peer may still be reading it
  epoch_end(sv, e, 2);
}

// ------------------------------------------------------------------ NVLS (multimem)
template <typename T> __device__ __forceinline__ uint4 multimem_ld_reduce(const void* mc);
template <> __device__ __forceinline__ uint4 multimem_ld_reduce<float>(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
template <> __device__ __forceinline__ uint4 multimem_ld_reduce<__nv_bfloat16>(const void* mc) {
  uint4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
               : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(void* mc, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

template <typename T>
__global__ void __launch_bounds__(AR_THREADS)
nvls_kernel(SymmView sv, size_t offset, size_t nvec_total, float scale) {
  constexpr int N = Pack<T>::N;
  const unsigned int e = epoch_begin(sv);
  group_barrier(sv, e + 1);
  const size_t per = (nvec_total + sv.world - 1) / sv.world;
  const size_t begin = min(nvec_total, per * sv.rank), end = min(nvec_total, begin + per);
  char* mc = static_cast<char*>(sv.mc) + offset;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t v0 = begin + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < end; v0 += stride * AR_UNROLL) {
    uint4 raw[AR_UNROLL];
#pragma unroll
    for (int u = 0; u < AR_UNROLL; ++u)
      if (v0 + u * stride < end) raw[u] = multimem_ld_reduce<T>(mc + (v0 + u * stride) * 16);
#pragma unroll
    for (int u = 0; u < AR_UNROLL; ++u) {
      const size_t v = v0 + u * stride;
      if (v >= end) continue;
      float f[N];
      Pack<T>::load(&raw[u], f);
#pragma unroll
      for (int i = 0; i < N; ++i) f[i] *= scale;
      Pack<T>::store(&raw[u], f);
      multimem_st(mc + v * 16, raw[u]);
    }
  }
  __threadfence_system();
  group_barrier(sv, e + 2);
  epoch_end(sv, e, 2);
}

// ------------------------------------------------------------------ fused reduce-scatter + SGD + parameter multicast
// One kernel per training step in place of {all-reduce, full-arena optimizer}: rank r owns vectors [r*per, (r+1)*per) of the
// flat arenas. After the entry barrier (every rank's gradients are complete) it pulls the sum of its slice of the gradient
// arena (NVLS: one multimem.ld_reduce per vector, the switch adds; P2P: one load per peer, added in rank order), applies
// 1/world and the SGD/momentum update to its slice of the parameters (momentum exists only on the owner), and stores the new
// parameters into EVERY replica's parameter arena (multimem.st, or one store per peer). The exit barrier makes all pushes
// visible before any rank's next forward pass reads them. Replicas stay bit-identical because each element is computed once.
// The last block to finish advances the step counter / first-step flag and logs the step's loss to pinned host memory.
template <int WORLD, bool NVLS>
__global__ void __launch_bounds__(AR_THREADS)
sgd_fused_kernel(SymmView sv, size_t g_off, size_t p_off, size_t m_off, size_t nvec_total, float inv_world, rtdc_sgd_hyper* hp,
                 uint32_t* step_counter, const float* loss, float* loss_log, uint32_t log_mask) {
  pdl_launch_dependents();
  pdl_wait();
  constexpr int U = NVLS ? AR_UNROLL : P2PUnroll<WORLD>::value;
  const unsigned int e = epoch_begin(sv);
  const rtdc_sgd_hyper h = *hp;
  group_barrier(sv, e + 1);
  const int world = NVLS ? sv.world : WORLD;
  const size_t per = (nvec_total + world - 1) / world;
  const size_t begin = min(nvec_total, per * sv.rank), end = min(nvec_total, begin + per);
  char* local = static_cast<char*>(sv.peers[sv.rank]);
  char* mc = static_cast<char*>(sv.mc);
  const bool use_m = h.momentum != 0.f;
  const size_t stride = static_cast<size_t>(gridDim.x) * blockDim.x;
  for (size_t v0 = begin + static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v0 < end; v0 += stride * U) {
    uint4 raw[U][WORLD];   // NVLS: one switch-reduced vector; P2P: one vector per peer, all in flight before the first add
    float4 pv[U], mv[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + u * stride;
      if (v >= end) continue;
      if (NVLS) {
        raw[u][0] = multimem_ld_reduce<float>(mc + g_off + v * 16);
      } else {
#pragma unroll
        for (int p = 0; p < WORLD; ++p) raw[u][p] = ld_peer16(static_cast<const char*>(sv.peers[p]) + g_off + v * 16);
      }
      pv[u] = *reinterpret_cast<const float4*>(local + p_off + v * 16);
      mv[u] = use_m ? *reinterpret_cast<const float4*>(local + m_off + v * 16) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const size_t v = v0 + u * stride;
      if (v >= end) continue;
      float gg[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int p = 0; p < WORLD; ++p) {   // fixed rank order (one pass for NVLS)
        float f[4];
        Pack<float>::load(&raw[u][p], f);
#pragma unroll
        for (int j = 0; j < 4; ++j) gg[j] += f[j];
      }
      float* pp = &pv[u].x;
      float* mm = &mv[u].x;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        float gr = gg[j] * inv_world * h.grad_scale + h.weight_decay * pp[j];
        if (use_m) {
          mm[j] = h.first_step ? gr : h.momentum * mm[j] + (1.f - h.dampening) * gr;
          gr = h.nesterov ? gr + h.momentum * mm[j] : mm[j];
        }
        pp[j] -= h.lr * gr;
      }
      if (use_m) *reinterpret_cast<float4*>(local + m_off + v * 16) = mv[u];
      const uint4 out = *reinterpret_cast<const uint4*>(&pv[u]);
      if (NVLS) {
        multimem_st(mc + p_off + v * 16, out);
      } else {
#pragma unroll
        for (int p = 0; p < WORLD; ++p) st_peer16(static_cast<char*>(sv.peers[(sv.rank + p) % WORLD]) + p_off + v * 16, out);
      }
    }
  }
  if (NVLS) __threadfence_system();
  group_barrier(sv, e + 2);
  epoch_end(sv, e, 2);
  if (threadIdx.x == 0) {
    __threadfence();
    const int ticket = atomicAdd(&hp->pad, 1);
    if (ticket == static_cast<int>(gridDim.x) - 1) {
      hp->first_step = 0;
      hp->pad = 0;
      if (step_counter != nullptr) {
        const uint32_t c = step_counter[0];
        if (loss_log != nullptr && loss != nullptr) loss_log[c & log_mask] = *loss;
        step_counter[0] = c + 1u;
      }
      __threadfence();
    }
  }
}

// Pull every other rank's slice of a sliced arena (the momentum of the fused step) into this rank's copy. No barrier: the
// caller guarantees the owners are not writing (they cannot start the next step before this rank joins its entry barrier).
__global__ void __launch_bounds__(AR_THREADS)
gather_slices_kernel(SymmView sv, size_t offset, size_t nvec_total) {
  const size_t per = (nvec_total + sv.world - 1) / sv.world;
  char* local = static_cast<char*>(sv.peers[sv.rank]) + offset;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec_total; v += static_cast<size_t>(gridDim.x) * blockDim.x) {
    const int owner = static_cast<int>(v / per);
    if (owner == sv.rank) continue;
    reinterpret_cast<uint4*>(local)[v] = ld_peer16(static_cast<const char*>(sv.peers[owner]) + offset + v * 16);
  }
}

// ------------------------------------------------------------------ broadcast / barrier
__global__ void __launch_bounds__(AR_THREADS)
broadcast_kernel(SymmView sv, size_t offset, size_t nvec, size_t tail_bytes, int root, void* dst) {
  const unsigned int e = epoch_begin(sv);
  group_barrier(sv, e + 1);
  const char* src = static_cast<const char*>(sv.peers[root]) + offset;
  for (size_t v = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x; v < nvec; v += static_cast<size_t>(gridDim.x) * blockDim.x)
    reinterpret_cast<uint4*>(dst)[v] = ld_peer16(src + v * 16);
  if (blockIdx.x == 0 && threadIdx.x < tail_bytes) {
    unsigned int b;
    asm volatile("ld.global.relaxed.sys.u8 %0, [%1];" : "=r"(b) : "l"(src + nvec * 16 + threadIdx.x) : "memory");
    static_cast<unsigned char*>(dst)[nvec * 16 + threadIdx.x] = static_cast<unsigned char>(b);
  }
  group_barrier(sv, e + 2);  // root may reuse the staging area afterwards
  epoch_end(sv, e, 2);
}

__global__ void barrier_kernel(SymmView sv) {
  const unsigned int e = epoch_begin(sv);
  group_barrier(sv, e + 1);
  epoch_end(sv, e, 1);
}

static int view_of(void* handle, SymmView* sv, rtdc_symm_info_t* info) {
  int rc = rtdc_symm_info(handle, info);
  if (rc != RTDC_OK) return rc;
  sv->peers = info->peer_table_dev; sv->pads = info->pad_table_dev; sv->epochs = info->epochs_dev;
  sv->mc = info->multicast_ptr; sv->rank = info->rank; sv->world = info->world;
  return RTDC_OK;
}

template <typename T>
static int launch_ar(const SymmView& sv, int world, int algo, size_t offset, size_t nvec, float scale, void* out_local, int blocks, cudaStream_t st) {
#define RTDC_AR_CASE(W)                                                                                           \
  case W:                                                                                                         \
    if (algo == 1) one_shot_kernel<T, W><<<blocks, AR_THREADS, 0, st>>>(sv, offset, nvec, scale, out_local);       \
    else two_shot_kernel<T, W><<<blocks, AR_THREADS, 0, st>>>(sv, offset, nvec, scale);                            \
    break;
  if (algo == 3) {
    nvls_kernel<T><<<blocks, AR_THREADS, 0, st>>>(sv, offset, nvec, scale);
  } else {
    switch (world) {
      RTDC_AR_CASE(1) RTDC_AR_CASE(2) RTDC_AR_CASE(3) RTDC_AR_CASE(4) RTDC_AR_CASE(5) RTDC_AR_CASE(6) RTDC_AR_CASE(7) RTDC_AR_CASE(8)
      default: return fail(RTDC_ERR_LIMIT, "all_reduce supports world sizes 1..8 (got %d)", world);
    }
  }
#undef RTDC_AR_CASE
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

}  // namespace rtdc

using namespace rtdc;

RTDC_API int rtdc_symm_all_reduce(void* handle, size_t offset, size_t count, int dtype, float scale, int algo,
                                  void* out_local, int max_blocks, void* stream) {
  SymmView sv;
  rtdc_symm_info_t info;
  int rc = view_of(handle, &sv, &info);
  if (rc != RTDC_OK) return rc;
  const size_t esize = dtype == 0 ? 4 : 2;
  if (dtype != 0 && dtype != 1) return fail(RTDC_ERR_ARG, "all_reduce: dtype must be 0 (fp32) or 1 (bf16)");
  if (offset % 16 != 0 || (count * esize) % 16 != 0)
    return fail(RTDC_ERR_ARG, "all_reduce: offset and byte length must be multiples of 16 (got %zu, %zu)", offset, count * esize);
  if (offset + count * esize > info.bytes) return fail(RTDC_ERR_LIMIT, "all_reduce: range exceeds the symmetric buffer (%zu bytes)", info.bytes);
  if (count == 0) return RTDC_OK;
  const size_t nvec = count * esize / 16;
  if (algo == 0) {
    const size_t bytes = count * esize;
    // measured on 8 B200 (tools/bench_allreduce.py, profiles/): the unrolled P2P two-shot wins up to tens of MB (2.68 MB: 24.8 vs
    // 25.9 us NVLS vs 40.3 us NCCL; 27 MiB: 88 vs 99 vs 117 us); the switch reduction wins once the message is bandwidth-bound
    // (98 MiB: 264 us NVLS vs 292 us two-shot vs 321 us NCCL)
    if (info.multicast_ptr && bytes >= (64u << 20) && info.world > 2) algo = 3;
    else algo = 2;
  }
  if (algo == 3 && !info.multicast_ptr) return fail(RTDC_ERR_ARG, "all_reduce: NVLS requested but multicast is not available");
  if (algo == 1) {
    if (!out_local) return fail(RTDC_ERR_ARG, "all_reduce: one-shot needs out_local");
    const char* lo = static_cast<const char*>(info.local_ptr);
    const char* o = static_cast<const char*>(out_local);
    if (o + count * esize > lo && o < lo + info.bytes) return fail(RTDC_ERR_ARG, "all_reduce: one-shot output must not alias the symmetric buffer");
  }
  // enough blocks to cover the owned slice, bounded by the signal-pad rows
  const size_t work = algo == 1 ? nvec : (nvec + info.world - 1) / info.world;
  int blocks = static_cast<int>((work + AR_THREADS * AR_UNROLL - 1) / (AR_THREADS * AR_UNROLL));
  // latency-bound sizes (the MLP's buckets) want few blocks (fewer barrier participants); bandwidth-bound ones want the links
  // saturated: ~1 block per SM
  int cap = max_blocks > 0 ? max_blocks : (count * esize >= (8u << 20) ? RTDC_SYMM_MAX_BLOCKS : 32);
  if (cap > RTDC_SYMM_MAX_BLOCKS) cap = RTDC_SYMM_MAX_BLOCKS;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  if (dtype == 0) return launch_ar<float>(sv, info.world, algo, offset, nvec, scale, out_local, blocks, st);
  return launch_ar<__nv_bfloat16>(sv, info.world, algo, offset, nvec, scale, out_local, blocks, st);
}

RTDC_API int rtdc_symm_broadcast(void* handle, size_t offset, size_t bytes, int root, void* dst_local, void* stream) {
  SymmView sv;
  rtdc_symm_info_t info;
  int rc = view_of(handle, &sv, &info);
  if (rc != RTDC_OK) return rc;
  if (root < 0 || root >= info.world) return fail(RTDC_ERR_ARG, "broadcast: bad root %d", root);
  if (offset % 16 != 0 || (reinterpret_cast<uintptr_t>(dst_local) & 15u)) return fail(RTDC_ERR_ARG, "broadcast: offset and destination must be 16-byte aligned");
  if (offset + bytes > info.bytes) return fail(RTDC_ERR_LIMIT, "broadcast: range exceeds the symmetric buffer");
  const size_t nvec = bytes / 16, tail = bytes % 16;
  int blocks = static_cast<int>((nvec + AR_THREADS * 4 - 1) / (AR_THREADS * 4));
  if (blocks > 32) blocks = 32;
  if (blocks < 1) blocks = 1;
  broadcast_kernel<<<blocks, AR_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(sv, offset, nvec, tail, root, dst_local);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API int rtdc_symm_barrier(void* handle, void* stream) {
  SymmView sv;
  rtdc_symm_info_t info;
  int rc = view_of(handle, &sv, &info);
  if (rc != RTDC_OK) return rc;
  barrier_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream)>>>(sv);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API int rtdc_symm_sgd_step(void* handle, size_t g_offset, size_t p_offset, size_t m_offset, size_t count, int algo,
                                rtdc_sgd_hyper* hyper_dev, uint32_t* step_counter_or_null, const float* loss_dev_or_null,
                                float* loss_log_host_or_null, uint32_t loss_log_slots, int max_blocks, void* stream) {
  SymmView sv;
  rtdc_symm_info_t info;
  int rc = view_of(handle, &sv, &info);
  if (rc != RTDC_OK) return rc;
  if (!hyper_dev) return fail(RTDC_ERR_ARG, "symm_sgd_step: null hyper-parameter pointer");
  if ((g_offset | p_offset | m_offset) % 16 != 0 || count % 4 != 0)
    return fail(RTDC_ERR_ARG, "symm_sgd_step: offsets must be multiples of 16 bytes and count a multiple of 4 elements");
  const size_t bytes = count * 4;
  if (g_offset + bytes > info.bytes || p_offset + bytes > info.bytes || m_offset + bytes > info.bytes)
    return fail(RTDC_ERR_LIMIT, "symm_sgd_step: an arena exceeds the symmetric buffer (%zu bytes)", info.bytes);
  if (loss_log_slots & (loss_log_slots - 1)) return fail(RTDC_ERR_ARG, "symm_sgd_step: loss_log_slots must be a power of two");
  if (count == 0) return RTDC_OK;
  // latency-bound at the MLP's 2.68 MB: peer loads + peer stores (24.3 us at 8 GPUs) beat the multimem path (28.9 us)
  if (algo == 0) algo = (info.multicast_ptr && info.world > 2 && bytes >= (64u << 20)) ? 3 : 2;
  if (algo == 3 && !info.multicast_ptr) return fail(RTDC_ERR_ARG, "symm_sgd_step: NVLS requested but multicast is not available");
  if (algo != 2 && algo != 3) return fail(RTDC_ERR_ARG, "symm_sgd_step: algo must be 0 (auto), 2 (P2P) or 3 (NVLS)");
  const size_t nvec = count / 4;
  const size_t work = (nvec + info.world - 1) / info.world;
  // Latency-bound at the MLP's size: every block is one more participant in both cross-GPU barriers. Measured at 2 GPUs on
  // the 2.68 MB arena (tools/bench_allreduce.py): 160 blocks 29.7 us, 64 blocks 19.7 us, 32 blocks 21.4 us, 16 blocks 25.6 us.
  int blocks = static_cast<int>((work + AR_THREADS * 2 - 1) / (AR_THREADS * 2));
  int cap = max_blocks > 0 ? max_blocks : 64;
  if (cap > RTDC_SYMM_MAX_BLOCKS) cap = RTDC_SYMM_MAX_BLOCKS;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  const float inv_world = 1.0f / static_cast<float>(info.world);
  const uint32_t mask = loss_log_slots ? loss_log_slots - 1 : 0;
  float* log = loss_log_slots ? loss_log_host_or_null : nullptr;
#define RTDC_SGD_CASE(W)                                                                                                  \
  case W:                                                                                                                 \
    sgd_fused_kernel<W, false><<<blocks, AR_THREADS, 0, st>>>(sv, g_offset, p_offset, m_offset, nvec, inv_world, hyper_dev, \
                                                              step_counter_or_null, loss_dev_or_null, log, mask);         \
    break;
  if (algo == 3) {
    sgd_fused_kernel<1, true><<<blocks, AR_THREADS, 0, st>>>(sv, g_offset, p_offset, m_offset, nvec, inv_world, hyper_dev,
                                                             step_counter_or_null, loss_dev_or_null, log, mask);
  } else {
    switch (info.world) {
      RTDC_SGD_CASE(1) RTDC_SGD_CASE(2) RTDC_SGD_CASE(3) RTDC_SGD_CASE(4) RTDC_SGD_CASE(5) RTDC_SGD_CASE(6) RTDC_SGD_CASE(7) RTDC_SGD_CASE(8)
      default: return fail(RTDC_ERR_LIMIT, "symm_sgd_step supports world sizes 1..8 (got %d)", info.world);
    }
  }
#undef RTDC_SGD_CASE
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}

RTDC_API int rtdc_symm_gather_slices(void* handle, size_t offset, size_t count, void* stream) {
  SymmView sv;
  rtdc_symm_info_t info;
  int rc = view_of(handle, &sv, &info);
  if (rc != RTDC_OK) return rc;
  if (offset % 16 != 0 || count % 4 != 0) return fail(RTDC_ERR_ARG, "gather_slices: offset must be a multiple of 16 bytes and count of 4 elements");
  if (offset + count * 4 > info.bytes) return fail(RTDC_ERR_LIMIT, "gather_slices: range exceeds the symmetric buffer");
  if (count == 0 || info.world == 1) return RTDC_OK;
  const size_t nvec = count / 4;
  int blocks = static_cast<int>((nvec + AR_THREADS * 4 - 1) / (AR_THREADS * 4));
  if (blocks > 148) blocks = 148;
  gather_slices_kernel<<<blocks, AR_THREADS, 0, static_cast<cudaStream_t>(stream)>>>(sv, offset, nvec);
  RTDC_CUDA_CHECK(cudaGetLastError());
  return RTDC_OK;
}
