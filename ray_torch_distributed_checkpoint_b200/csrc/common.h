// Host-side helpers shared by every translation unit of librtdc_b200.so.
#pragma once
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdarg.h>
#include <stdio.h>
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

inline int ceil_div(int a, int b) { return (a + b - 1) / b; }
inline int round_up(int a, int b) { return ceil_div(a, b) * b; }

}  // namespace rtdc
