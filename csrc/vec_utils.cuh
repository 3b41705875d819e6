// vec_utils.cuh — 4-wide typed loads/stores (fp32 / bf16 <-> float registers), warp/block reductions and a
// counter-based Philox4x32-10 generator used for dropout masks that backward can regenerate.
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>

namespace b200 {

// programmatic dependent launch (see launch_pdl in common.h); both are no-ops for ordinary launches
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }


struct f4 {
  float x, y, z, w;
};

template <typename T>
__device__ __forceinline__ f4 load4(const T* p);
template <>
__device__ __forceinline__ f4 load4<float>(const float* p) {
  const float4 v = *reinterpret_cast<const float4*>(p);
  return {v.x, v.y, v.z, v.w};
}
template <>
__device__ __forceinline__ f4 load4<__nv_bfloat16>(const __nv_bfloat16* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x);
  const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
  return {__low2float(a), __high2float(a), __low2float(b), __high2float(b)};
}
template <typename T>
__device__ __forceinline__ void store4(T* p, const f4& v);
template <>
__device__ __forceinline__ void store4<float>(float* p, const f4& v) {
  *reinterpret_cast<float4*>(p) = make_float4(v.x, v.y, v.z, v.w);
}
template <>
__device__ __forceinline__ void store4<__nv_bfloat16>(__nv_bfloat16* p, const f4& v) {
  __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y);
  __nv_bfloat162 b = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<uint32_t*>(&a);
  u.y = *reinterpret_cast<uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}
template <typename T>
__device__ __forceinline__ float load1(const T* p);
template <>
__device__ __forceinline__ float load1<float>(const float* p) { return *p; }
template <>
__device__ __forceinline__ float load1<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }
template <typename T>
__device__ __forceinline__ void store1(T* p, float v);
template <>
__device__ __forceinline__ void store1<float>(float* p, float v) { *p = v; }
template <>
__device__ __forceinline__ void store1<__nv_bfloat16>(__nv_bfloat16* p, float v) { *p = __float2bfloat16_rn(v); }

// ---- cp.async (LDGSTS): per-thread asynchronous global -> shared copies, completion tracked per thread in commit
// groups.  A thread that only reads back what it copied itself needs no barrier: wait_group makes its own copies
// visible to it.  Used as a register-free prefetch buffer by the streaming LayerNorm kernels.
template <int BYTES>
__device__ __forceinline__ void cp_async(void* smem_dst, const void* gsrc) {
  const uint32_t d = (uint32_t)__cvta_generic_to_shared(smem_dst);
  if constexpr (BYTES == 16) {
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gsrc) : "memory");
  } else {
    asm volatile("cp.async.ca.shared.global [%0], [%1], %2;" ::"r"(d), "l"(gsrc), "n"(BYTES) : "memory");
  }
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

// storage form of four consecutive elements (what load4 reads) and its conversion to registers
template <typename T> struct Raw4;
template <> struct Raw4<float> {
  using type = float4;
  static __device__ __forceinline__ f4 cvt(const float4& v) { return {v.x, v.y, v.z, v.w}; }
};
template <> struct Raw4<__nv_bfloat16> {
  using type = uint2;
  static __device__ __forceinline__ f4 cvt(const uint2& u) {
    const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x);
    const __nv_bfloat162 b = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
    return {__low2float(a), __high2float(a), __low2float(b), __high2float(b)};
  }
};

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// ---- Philox4x32-10 ---------------------------------------------------------------------------------
struct Philox {
  uint32_t k0, k1;
  __device__ __forceinline__ Philox(unsigned long long seed) : k0((uint32_t)seed), k1((uint32_t)(seed >> 32)) {}
  // 4 x 32 random bits for counter `ctr` (128-bit counter, upper half = stream id)
  __device__ __forceinline__ uint4 operator()(unsigned long long ctr, unsigned long long stream) const {
    uint32_t c0 = (uint32_t)ctr, c1 = (uint32_t)(ctr >> 32), c2 = (uint32_t)stream, c3 = (uint32_t)(stream >> 32);
    uint32_t a = k0, b = k1;
#pragma unroll
    for (int r = 0; r < 10; ++r) {
      const uint32_t hi0 = __umulhi(0xD2511F53u, c0), lo0 = 0xD2511F53u * c0;
      const uint32_t hi1 = __umulhi(0xCD9E8D57u, c2), lo1 = 0xCD9E8D57u * c2;
      const uint32_t n0 = hi1 ^ c1 ^ a, n1 = lo1, n2 = hi0 ^ c3 ^ b, n3 = lo0;
      c0 = n0; c1 = n1; c2 = n2; c3 = n3;
      a += 0x9E3779B9u;
      b += 0xBB67AE85u;
    }
    return make_uint4(c0, c1, c2, c3);
  }
};

// keep-mask scale for 4 consecutive elements starting at element index `idx` (idx % 4 == 0):
// returns 1/(1-p) where kept, 0 where dropped.  p_thresh = p * 2^32.
__device__ __forceinline__ f4 dropout_scale4(const Philox& ph, unsigned long long idx, unsigned long long stream,
                                             uint32_t p_thresh, float inv_keep) {
  const uint4 r = ph(idx >> 2, stream);
  return {r.x >= p_thresh ? inv_keep : 0.f, r.y >= p_thresh ? inv_keep : 0.f, r.z >= p_thresh ? inv_keep : 0.f,
          r.w >= p_thresh ? inv_keep : 0.f};
}

__host__ __device__ inline uint32_t dropout_threshold(float p) {
  double t = (double)p * 4294967296.0;
  if (t < 0) t = 0;
  if (t > 4294967295.0) t = 4294967295.0;
  return (uint32_t)t;
}

}  // namespace b200
