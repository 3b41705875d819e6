// common.h — host-side helpers shared by every translation unit of libb200seq.so.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdio>

namespace b200 {

enum : int { B200_ERR_ARG = -1, B200_ERR_UNSUPPORTED = -2, B200_ERR_DRIVER = -3, B200_ERR_STATE = -4 };

extern thread_local char g_err[512];
extern std::atomic<long long> g_launches;

inline int fail(int code, const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
  return code;
}

inline int cuda_fail(cudaError_t e, const char* what) {
  snprintf(g_err, sizeof(g_err), "%s: %s", what, cudaGetErrorString(e));
  return (int)e;
}

#define B200_CUDA(expr)                                        \
  do {                                                         \
    cudaError_t _e = (expr);                                   \
    if (_e != cudaSuccess) return ::b200::cuda_fail(_e, #expr); \
  } while (0)

// Call right after a <<<>>> launch.
#define B200_LAUNCHED(name)                                          \
  do {                                                               \
    ::b200::g_launches.fetch_add(1, std::memory_order_relaxed);      \
    cudaError_t _e = cudaGetLastError();                             \
    if (_e != cudaSuccess) return ::b200::cuda_fail(_e, name);       \
  } while (0)

inline int num_sms() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev);
    if (n <= 0) n = 148;
  }
  return n;
}

}  // namespace b200
