// common.h — host-side helpers shared by every translation unit of libb200seq.so.
#pragma once
#include <cuda_runtime.h>
#include <atomic>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>

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

// Programmatic dependent launch: a kernel launched through launch_pdl may be scheduled while its predecessor on
// the stream is still draining; it must execute pdl_wait() (vec_utils.cuh) before touching memory the predecessor
// produces.  Kernels call pdl_trigger() first thing so that THEIR successor can be scheduled early.  Hides launch
// latency and per-kernel prologues (barrier init, TMEM allocation, descriptor prefetch) behind the previous tail.
// Opt-in (B200_PDL=1): measured on B200 it removes ~0.7 us per kernel boundary inside a captured graph (attention
// forward 25.6 -> 25.0 us back to back) but the whole training step does not move (8.616 vs 8.598 ms), so the
// default stays with ordinary stream-ordered launches.
inline bool pdl_enabled() {
  static const bool on = [] { const char* e = getenv("B200_PDL"); return e && e[0] == '1'; }();
  return on;
}

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                              int cluster_x, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[2];
  unsigned n = 0;
  if (pdl_enabled()) {
    attr[n].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[n].val.programmaticStreamSerializationAllowed = 1;
    ++n;
  }
  if (cluster_x > 1) {
    attr[n].id = cudaLaunchAttributeClusterDimension;
    attr[n].val.clusterDim.x = cluster_x;
    attr[n].val.clusterDim.y = 1;
    attr[n].val.clusterDim.z = 1;
    ++n;
  }
  cfg.attrs = attr;
  cfg.numAttrs = n;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

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

// SMs the persistent compute kernels (GEMM) may fill.  Data-parallel runs can leave a few SMs to the gradient
// reduction kernels that overlap backward (B200_COMPUTE_SMS, set by the reducer or by hand); default: all.
inline std::atomic<int>& compute_sm_limit() {
  static std::atomic<int> v{0};
  return v;
}
inline int compute_sms() {
  int lim = compute_sm_limit().load(std::memory_order_relaxed);
  if (lim == 0) {
    const char* e = getenv("B200_COMPUTE_SMS");
    lim = e ? atoi(e) : -1;
    if (lim == 0) lim = -1;
    compute_sm_limit().store(lim, std::memory_order_relaxed);
  }
  const int n = num_sms();
  return (lim > 0 && lim < n) ? lim : n;
}

}  // namespace b200
