// optim_kernels.cu — flat-arena optimizer step for sm_100a.
//
// The reference scales, clips and updates with ~2.5k tiny launches and three host syncs per step
// (hetseq/optim.py:59-70,162-231; hetseq/controller.py:338-363).  Here the whole parameter set lives in
// one fp32 arena, so a step is: (1) one deterministic sum-of-squares reduction, (2) one fused kernel that
// derives scale = mult / sample_size and the clip coefficient ON THE DEVICE, applies Adam (reference
// math, decoupled weight decay, eps outside the sqrt), optionally refreshes the bf16 working copy and
// zeroes the gradient for the next step.  No host round trip anywhere.
#include "common.h"
#include "vec_utils.cuh"
#include "optim_math.cuh"
#include "../include/b200seq.h"

namespace b200 {

constexpr int kSumsqThreads = 512;
constexpr int kSumsqMaxBlocks = 1184;  // 8 per SM on a 148-SM part

// ---- sum of squares: per-block partials into scratch, last block folds them in a fixed order ----------
// scratch layout: [0, kSumsqMaxBlocks) float partials, then one uint32 arrival counter (must start at 0;
// the kernel leaves it at 0 again).
__global__ void __launch_bounds__(kSumsqThreads)
sumsq_kernel(const float* __restrict__ g, long long n, float* __restrict__ out, float* __restrict__ scratch,
             int accumulate) {
  __shared__ float warp_part[kSumsqThreads / 32];
  __shared__ bool is_last;
  float acc = 0.f;
  const long long n4 = n >> 2;
  const float4* g4 = reinterpret_cast<const float4*>(g);
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = __ldg(g4 + i);
    acc += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const float v = g[(n4 << 2) + threadIdx.x];
    acc += v * v;
  }
  acc = warp_sum(acc);
  if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < kSumsqThreads / 32 ? warp_part[threadIdx.x] : 0.f;
    v = warp_sum(v);
    if (threadIdx.x == 0) {
      scratch[blockIdx.x] = v;
      __threadfence();
      unsigned int* counter = reinterpret_cast<unsigned int*>(scratch + kSumsqMaxBlocks);
      const unsigned int prev = atomicAdd(counter, 1u);
      is_last = (prev == gridDim.x - 1);
    }
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    // fixed-order fold: thread t sums partials t, t+512, ... then the block tree-reduces
    float v = 0.f;
    for (int i = threadIdx.x; i < (int)gridDim.x; i += blockDim.x) v += __ldcg(scratch + i);
    v = warp_sum(v);
    if ((threadIdx.x & 31) == 0) warp_part[threadIdx.x >> 5] = v;
    __syncthreads();
    if (threadIdx.x < 32) {
      float t = threadIdx.x < kSumsqThreads / 32 ? warp_part[threadIdx.x] : 0.f;
      t = warp_sum(t);
      if (threadIdx.x == 0) {
        out[0] = accumulate ? out[0] + t : t;
        *reinterpret_cast<unsigned int*>(scratch + kSumsqMaxBlocks) = 0u;
      }
    }
  }
}

static int sumsq_grid(long long n) {
  long long blocks = (n / 4 + kSumsqThreads * 4 - 1) / (kSumsqThreads * 4);
  if (blocks < 1) blocks = 1;
  if (blocks > kSumsqMaxBlocks) blocks = kSumsqMaxBlocks;
  return (int)blocks;
}

__global__ void __launch_bounds__(256)
adam_flat_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ m, float* __restrict__ v,
                 __nv_bfloat16* __restrict__ p16, long long n, AdamArgsDev a) {
  float decay = 1.f - a.weight_decay * a.lr;
  if (a.hyper_dev != nullptr) {            // values that change every step come from device memory (graph replay)
    a.lr = a.hyper_dev[0];
    a.step_size = a.hyper_dev[1];
    decay = a.hyper_dev[2];
  }
  const StepScalars sc = derive_scalars(a.mult, a.sample_size, a.sample_size_f32, a.norm_sq, a.max_norm);
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.stats_out != nullptr) {
    a.stats_out[0] = sc.gnorm;
    a.stats_out[1] = sc.scale;
    a.stats_out[2] = sc.clipped;
  }
  const float one_m_b1 = 1.f - a.beta1, one_m_b2 = 1.f - a.beta2;
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    float4 pv = reinterpret_cast<float4*>(p)[i];
    const float4 gv = reinterpret_cast<const float4*>(g)[i];
    float4 mv = reinterpret_cast<float4*>(m)[i];
    float4 vv = reinterpret_cast<float4*>(v)[i];
    float* pp = &pv.x; const float* gp = &gv.x; float* mp = &mv.x; float* vp = &vv.x;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float gs = gp[e] * sc.scale;
      mp[e] = a.beta1 * mp[e] + one_m_b1 * gs;
      vp[e] = a.beta2 * vp[e] + one_m_b2 * gs * gs;
      const float denom = sqrtf(vp[e]) + a.eps;
      float w = pp[e];
      if (a.weight_decay != 0.f) w *= decay;
      pp[e] = w - a.step_size * (mp[e] / denom);
    }
    reinterpret_cast<float4*>(p)[i] = pv;
    reinterpret_cast<float4*>(m)[i] = mv;
    reinterpret_cast<float4*>(v)[i] = vv;
    if (a.zero_grad) reinterpret_cast<float4*>(g)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p16 != nullptr) store4<__nv_bfloat16>(p16 + (i << 2), f4{pv.x, pv.y, pv.z, pv.w});
  }
  // tail (n not a multiple of 4)
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long long i = (n4 << 2) + threadIdx.x;
    const float gs = g[i] * sc.scale;
    const float mm = a.beta1 * m[i] + one_m_b1 * gs;
    const float vv = a.beta2 * v[i] + one_m_b2 * gs * gs;
    float w = p[i];
    if (a.weight_decay != 0.f) w *= decay;
    w -= a.step_size * (mm / (sqrtf(vv) + a.eps));
    p[i] = w; m[i] = mm; v[i] = vv;
    if (a.zero_grad) g[i] = 0.f;
    if (p16 != nullptr) p16[i] = __float2bfloat16_rn(w);
  }
}

struct AdadeltaArgsDev {
  float lr, rho, eps, weight_decay, mult, max_norm;
  const float* norm_sq;
  const double* sample_size;
  float* stats_out;
  int zero_grad;
  const float* sample_size_f32;
  const float* hyper_dev;
};

// reference math: hetseq/optim.py:293-302
__global__ void __launch_bounds__(256)
adadelta_flat_kernel(float* __restrict__ p, float* __restrict__ g, float* __restrict__ sq, float* __restrict__ acc,
                     __nv_bfloat16* __restrict__ p16, long long n, AdadeltaArgsDev a) {
  if (a.hyper_dev != nullptr) a.lr = a.hyper_dev[0];
  const StepScalars sc = derive_scalars(a.mult, a.sample_size, a.sample_size_f32, a.norm_sq, a.max_norm);
  if (blockIdx.x == 0 && threadIdx.x == 0 && a.stats_out != nullptr) {
    a.stats_out[0] = sc.gnorm;
    a.stats_out[1] = sc.scale;
    a.stats_out[2] = sc.clipped;
  }
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    float w = p[i];
    float gs = g[i] * sc.scale;
    if (a.weight_decay != 0.f) gs += a.weight_decay * w;
    const float s = a.rho * sq[i] + (1.f - a.rho) * gs * gs;
    const float stdv = sqrtf(s + a.eps);
    const float delta = sqrtf(acc[i] + a.eps) / stdv * gs;
    w -= a.lr * delta;
    p[i] = w;
    sq[i] = s;
    acc[i] = a.rho * acc[i] + (1.f - a.rho) * delta * delta;
    if (a.zero_grad) g[i] = 0.f;
    if (p16 != nullptr) p16[i] = __float2bfloat16_rn(w);
  }
}

__global__ void __launch_bounds__(256) scale_kernel(float* __restrict__ x, long long n, float c, const float* c_dev) {
  const float s = c_dev ? c * (*c_dev) : c;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) x[i] *= s;
}

__global__ void __launch_bounds__(256)
cast_f32_bf16_kernel(const float* __restrict__ src, __nv_bfloat16* __restrict__ dst, long long n) {
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(src)[i];
    store4<__nv_bfloat16>(dst + (i << 2), f4{v.x, v.y, v.z, v.w});
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long long i = (n4 << 2) + threadIdx.x;
    dst[i] = __float2bfloat16_rn(src[i]);
  }
}

// 3xTF32 operand split: hi = x with the 13 mantissa bits a TF32 operand drops cleared, lo = (x - hi) cut the same
// way.  hi and lo are exactly representable TF32 inputs; hi*hi' + hi*lo' + lo*hi' recovers the fp32 product to ~2^-22.
__global__ void __launch_bounds__(256)
split_tf32_kernel(const float* __restrict__ x, float* __restrict__ hi, float* __restrict__ lo, long long n) {
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const float4 v = reinterpret_cast<const float4*>(x)[i];
    float4 h, l;
    const float* vp = &v.x; float* hp = &h.x; float* lp = &l.x;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      hp[e] = __uint_as_float(__float_as_uint(vp[e]) & 0xFFFFE000u);
      lp[e] = __uint_as_float(__float_as_uint(vp[e] - hp[e]) & 0xFFFFE000u);
    }
    reinterpret_cast<float4*>(hi)[i] = h;
    reinterpret_cast<float4*>(lo)[i] = l;
  }
  if (blockIdx.x == 0 && threadIdx.x < (n & 3)) {
    const long long i = (n4 << 2) + threadIdx.x;
    const float h = __uint_as_float(__float_as_uint(x[i]) & 0xFFFFE000u);
    hi[i] = h;
    lo[i] = __uint_as_float(__float_as_uint(x[i] - h) & 0xFFFFE000u);
  }
}

static int elementwise_grid(long long work_items, int threads) {
  long long blocks = (work_items + threads - 1) / threads;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  return (int)blocks;
}

}  // namespace b200

using namespace b200;

extern "C" long long b200_sumsq_scratch_bytes(void) { return (long long)(kSumsqMaxBlocks + 1) * 4; }

extern "C" int b200_sumsq_f32(const float* g, long long n, float* out, void* scratch, long long scratch_bytes,
                              int accumulate, void* stream) {
  if (!g || !out || !scratch) return fail(B200_ERR_ARG, "sumsq: null pointer");
  if (scratch_bytes < b200_sumsq_scratch_bytes()) return fail(B200_ERR_ARG, "sumsq: scratch too small");
  if ((reinterpret_cast<uintptr_t>(g) & 15) != 0) return fail(B200_ERR_ARG, "sumsq: input must be 16-byte aligned");
  sumsq_kernel<<<sumsq_grid(n), kSumsqThreads, 0, (cudaStream_t)stream>>>(g, n, out, (float*)scratch, accumulate);
  B200_LAUNCHED("sumsq_kernel");
  return 0;
}

extern "C" int b200_adam_flat(float* p, float* g, float* m, float* v, void* p_bf16, long long n,
                              const b200_adam_args* a, void* stream) {
  if (!p || !g || !m || !v || !a) return fail(B200_ERR_ARG, "adam: null pointer");
  if (((reinterpret_cast<uintptr_t>(p) | reinterpret_cast<uintptr_t>(g) | reinterpret_cast<uintptr_t>(m) |
        reinterpret_cast<uintptr_t>(v)) & 15) != 0)
    return fail(B200_ERR_ARG, "adam: buffers must be 16-byte aligned");
  AdamArgsDev d;
  fill_adam_args(a, d);
  adam_flat_kernel<<<elementwise_grid(n / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>(p, g, m, v, (__nv_bfloat16*)p_bf16, n, d);
  B200_LAUNCHED("adam_flat_kernel");
  return 0;
}

extern "C" int b200_adadelta_flat(float* p, float* g, float* sq, float* acc, void* p_bf16, long long n,
                                  const b200_adadelta_args* a, void* stream) {
  if (!p || !g || !sq || !acc || !a) return fail(B200_ERR_ARG, "adadelta: null pointer");
  AdadeltaArgsDev d;
  d.lr = a->lr; d.rho = a->rho; d.eps = a->eps; d.weight_decay = a->weight_decay;
  d.mult = a->grad_scale_mult; d.max_norm = a->max_norm;
  d.norm_sq = a->norm_sq; d.sample_size = a->sample_size; d.stats_out = a->stats_out; d.zero_grad = a->zero_grad;
  d.sample_size_f32 = a->sample_size_f32; d.hyper_dev = a->hyper_dev;
  adadelta_flat_kernel<<<elementwise_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(p, g, sq, acc, (__nv_bfloat16*)p_bf16, n, d);
  B200_LAUNCHED("adadelta_flat_kernel");
  return 0;
}

extern "C" int b200_scale_f32(float* x, long long n, float c, const float* c_dev, void* stream) {
  if (!x) return fail(B200_ERR_ARG, "scale: null pointer");
  scale_kernel<<<elementwise_grid(n, 256), 256, 0, (cudaStream_t)stream>>>(x, n, c, c_dev);
  B200_LAUNCHED("scale_kernel");
  return 0;
}

extern "C" int b200_cast_f32_to_bf16(const float* src, void* dst, long long n, void* stream) {
  if (!src || !dst) return fail(B200_ERR_ARG, "cast: null pointer");
  cast_f32_bf16_kernel<<<elementwise_grid(n / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>(src, (__nv_bfloat16*)dst, n);
  B200_LAUNCHED("cast_f32_bf16_kernel");
  return 0;
}

extern "C" int b200_split_tf32(const float* x, float* hi, float* lo, long long n, void* stream) {
  if (!x || !hi || !lo) return fail(B200_ERR_ARG, "split_tf32: null pointer");
  if (((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(hi) | reinterpret_cast<uintptr_t>(lo)) & 15) != 0)
    return fail(B200_ERR_ARG, "split_tf32: buffers must be 16-byte aligned");
  split_tf32_kernel<<<elementwise_grid(n / 4 + 1, 256), 256, 0, (cudaStream_t)stream>>>(x, hi, lo, n);
  B200_LAUNCHED("split_tf32_kernel");
  return 0;
}
