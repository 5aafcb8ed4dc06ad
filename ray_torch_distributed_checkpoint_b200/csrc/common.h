// Host-side helpers shared by every translation unit of librtdc_b200.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#define RTDC_API extern "C" __attribute__((visibility("default")))

namespace rtdc {

// Last error message of the calling thread; read through rtdc_last_error().
inline char* error_buffer() {
  static thread_local char buf[512] = {0};
  return buf;
}
inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(error_buffer(), 512, fmt, ap);
  va_end(ap);
  return code;
}

enum : int {
  RTDC_OK = 0,
  RTDC_ERR_ARG = 1,      // caller passed something the kernel cannot run on
  RTDC_ERR_CUDA = 2,     // a CUDA runtime/driver call failed
  RTDC_ERR_LIMIT = 3,    // a size exceeds a documented limit
  RTDC_ERR_SYSTEM = 4,   // OS-level failure (sockets, fds, fork/exec)
};

#define RTDC_CUDA_CHECK(expr)                                                                      \
  do {                                                                                             \
    cudaError_t _e = (expr);                                                                       \
    if (_e != cudaSuccess)                                                                         \
      return ::rtdc::fail(::rtdc::RTDC_ERR_CUDA, "%s failed: %s (%s:%d)", #expr, cudaGetErrorString(_e), \
                          __FILE__, __LINE__);                                                     \
  } while (0)

// Launch with the programmatic-dependent-launch attribute (see ptx.cuh: pdl_wait) unless RTDC_PDL=0. Only for kernels that
// call pdl_wait() before their first access to global memory another kernel may have written (or still reads).
inline bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("RTDC_PDL"); return !(e && e[0] == '0'); }();
  return on;
}
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args&&... args) {
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int round_up(int a, int b) { return ceil_div(a, b) * b; }

}  // namespace rtdc
