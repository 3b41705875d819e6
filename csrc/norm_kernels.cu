// norm_kernels.cu — fused  y = post_dropout( LayerNorm( pre_dropout(x + bias) + residual ) )  forward and
// backward for sm_100a.
//
// Replaces the reference's unfused BertLayerNorm (mean, sub, pow, mean, add, sqrt, div, mul, add = 9
// kernels, hetseq/bert_modeling.py:276-289) together with the dropout + residual add that precede it
// (bert_modeling.py:387-391,423-427) and the dropout that follows the embedding LayerNorm (:306-320).
// TF-style statistics: biased variance, eps inside the sqrt, computed in fp32 regardless of storage type.
//
// Dropout masks are never stored: they are regenerated from (seed, call-site stream id, element index)
// with Philox4x32-10 in backward.  One warp owns one row; a lane holds NV float4 vectors of it.
#include "common.h"
#include "vec_utils.cuh"
#include "../include/b200seq.h"

namespace b200 {

constexpr int kLnWarps = 8;
constexpr int kLnThreads = kLnWarps * 32;
constexpr int kLnBwdMaxBlocks = 296;   // 2 per SM
constexpr int kLnBwdMaxBlocks3 = 444;  // 3 per SM (ln_bwd2_kernel<..., 3>)

template <typename T>
struct LnCommon {
  const T* x; long long ldx;
  const T* bias;
  const T* residual; long long ldr;
  const T* gamma; const T* beta;
  float eps, p_pre, p_post;
  const unsigned long long* seed_dev;
  unsigned long long stream_pre, stream_post;
  int rows, cols;
};

template <typename T, int NV>
__device__ __forceinline__ void ln_load_row(const LnCommon<T>& a, int row, int lane, const Philox& ph, uint32_t thr_pre,
                                            float inv_keep_pre, f4 (&v)[NV], f4 (&mask)[NV]) {
#pragma unroll
  for (int k = 0; k < NV; ++k) {
    const int c = (k * 32 + lane) * 4;
    v[k] = {0.f, 0.f, 0.f, 0.f};
    mask[k] = {1.f, 1.f, 1.f, 1.f};
    if (c < a.cols) {
      f4 t = load4<T>(a.x + (long long)row * a.ldx + c);
      if (a.bias) {
        const f4 b = load4<T>(a.bias + c);
        t.x += b.x; t.y += b.y; t.z += b.z; t.w += b.w;
      }
      if (a.p_pre > 0.f) {
        mask[k] = dropout_scale4(ph, (unsigned long long)row * a.cols + c, a.stream_pre, thr_pre, inv_keep_pre);
        t.x *= mask[k].x; t.y *= mask[k].y; t.z *= mask[k].z; t.w *= mask[k].w;
      }
      if (a.residual) {
        const f4 r = load4<T>(a.residual + (long long)row * a.ldr + c);
        t.x += r.x; t.y += r.y; t.z += r.z; t.w += r.w;
      }
      v[k] = t;
    }
  }
}

// PIPE: the rows of a warp stream through shared memory with cp.async, one row ahead: every thread copies the
// vectors it will itself consume (no barrier needed), so the loads of row i+1 are in flight while row i is being
// normalised and stored.  Without it all warps of the (single-wave) grid load, then compute, then store in step
// and the three phases add up: 12.8 us for 4096 x 768 fp32 against 5.8 us of HBM time.
template <typename T, int NV, bool PIPE>
__global__ void __launch_bounds__(kLnThreads, PIPE ? 2 : 4)
ln_fwd_kernel(const LnCommon<T> a, T* __restrict__ y, long long ldy, float* __restrict__ mean_out,
              float* __restrict__ rstd_out) {
  pdl_trigger();
  pdl_wait();
  using R = typename Raw4<T>::type;
  extern __shared__ __align__(16) unsigned char ln_pipe_smem[];   // PIPE: R [2 stages][2 tensors][NV][kLnThreads]
  R* const buf = reinterpret_cast<R*>(ln_pipe_smem);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned long long seed = a.seed_dev ? *a.seed_dev : 0ull;
  const Philox ph(seed);
  const uint32_t thr_pre = dropout_threshold(a.p_pre), thr_post = dropout_threshold(a.p_post);
  const float ikp = a.p_pre > 0.f ? 1.f / (1.f - a.p_pre) : 1.f;
  const float ikq = a.p_post > 0.f ? 1.f / (1.f - a.p_post) : 1.f;
  const float inv_cols = 1.f / (float)a.cols;
  const int stride = gridDim.x * kLnWarps;
  int row = blockIdx.x * kLnWarps + warp;
  auto issue = [&](int r, int stage) {
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * 4;
      if (c < a.cols) {
        cp_async<sizeof(R)>(buf + ((stage * 2 + 0) * NV + k) * kLnThreads + threadIdx.x, a.x + (long long)r * a.ldx + c);
        if (a.residual)
          cp_async<sizeof(R)>(buf + ((stage * 2 + 1) * NV + k) * kLnThreads + threadIdx.x, a.residual + (long long)r * a.ldr + c);
      }
    }
  };
  if constexpr (PIPE) {
    if (row < a.rows) issue(row, 0);
    cp_async_commit();
  }
  for (int it = 0; row < a.rows; row += stride, ++it) {
    f4 v[NV];
    if constexpr (PIPE) {
      if (row + stride < a.rows) issue(row + stride, (it + 1) & 1);
      cp_async_commit();                    // one group per iteration (possibly empty): the count below stays uniform
      cp_async_wait<1>();                   // everything but the newest group has landed: this row is in `it & 1`
      const int st = it & 1;
#pragma unroll
      for (int k = 0; k < NV; ++k) {
        const int c = (k * 32 + lane) * 4;
        v[k] = {0.f, 0.f, 0.f, 0.f};
        if (c < a.cols) {
          f4 t = Raw4<T>::cvt(buf[((st * 2 + 0) * NV + k) * kLnThreads + threadIdx.x]);
          if (a.bias) {
            const f4 b = load4<T>(a.bias + c);
            t.x += b.x; t.y += b.y; t.z += b.z; t.w += b.w;
          }
          if (a.p_pre > 0.f) {
            const f4 m = dropout_scale4(ph, (unsigned long long)row * a.cols + c, a.stream_pre, thr_pre, ikp);
            t.x *= m.x; t.y *= m.y; t.z *= m.z; t.w *= m.w;
          }
          if (a.residual) {
            const f4 r = Raw4<T>::cvt(buf[((st * 2 + 1) * NV + k) * kLnThreads + threadIdx.x]);
            t.x += r.x; t.y += r.y; t.z += r.z; t.w += r.w;
          }
          v[k] = t;
        }
      }
    } else {
      f4 mask[NV];
      ln_load_row<T, NV>(a, row, lane, ph, thr_pre, ikp, v, mask);
    }
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) s += v[k].x + v[k].y + v[k].z + v[k].w;
    const float mean = warp_sum(s) * inv_cols;
    float q = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * 4;
      if (c < a.cols) {
        const float d0 = v[k].x - mean, d1 = v[k].y - mean, d2 = v[k].z - mean, d3 = v[k].w - mean;
        q += d0 * d0 + d1 * d1 + d2 * d2 + d3 * d3;
      }
    }
    const float rstd = rsqrtf(warp_sum(q) * inv_cols + a.eps);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * 4;
      if (c < a.cols) {
        const f4 g = load4<T>(a.gamma + c), b = load4<T>(a.beta + c);
        f4 o = {(v[k].x - mean) * rstd * g.x + b.x, (v[k].y - mean) * rstd * g.y + b.y,
                (v[k].z - mean) * rstd * g.z + b.z, (v[k].w - mean) * rstd * g.w + b.w};
        if (a.p_post > 0.f) {
          const f4 m = dropout_scale4(ph, (unsigned long long)row * a.cols + c, a.stream_post, thr_post, ikq);
          o.x *= m.x; o.y *= m.y; o.z *= m.z; o.w *= m.w;
        }
        store4<T>(y + (long long)row * ldy + c, o);
      }
    }
    if (lane == 0) {
      mean_out[row] = mean;
      rstd_out[row] = rstd;
    }
  }
}

// Backward.  Per-thread register accumulators for d(gamma), d(beta), d(bias) over the rows this warp owns;
// block partials go to scratch[block][3][cols]; ln_bwd_finalize folds them (fixed order) into the fp32
// parameter-gradient arena with +=.
template <typename T, int NV>
__global__ void __launch_bounds__(kLnThreads)
ln_bwd_kernel(const LnCommon<T> a, const T* __restrict__ dy, long long lddy, const float* __restrict__ mean_in,
              const float* __restrict__ rstd_in, T* __restrict__ ds, long long ldds, T* __restrict__ dz, long long lddz,
              float* __restrict__ scratch, int want_dbias) {
  extern __shared__ float red[];  // [kLnWarps][cols]
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const unsigned long long seed = a.seed_dev ? *a.seed_dev : 0ull;
  const Philox ph(seed);
  const uint32_t thr_pre = dropout_threshold(a.p_pre), thr_post = dropout_threshold(a.p_post);
  const float ikp = a.p_pre > 0.f ? 1.f / (1.f - a.p_pre) : 1.f;
  const float ikq = a.p_post > 0.f ? 1.f / (1.f - a.p_post) : 1.f;
  const float inv_cols = 1.f / (float)a.cols;
  f4 acc_g[NV], acc_b[NV], acc_bias[NV];
#pragma unroll
  for (int k = 0; k < NV; ++k) acc_g[k] = acc_b[k] = acc_bias[k] = {0.f, 0.f, 0.f, 0.f};

  for (int row = blockIdx.x * kLnWarps + warp; row < a.rows; row += gridDim.x * kLnWarps) {
    f4 v[NV], mask[NV], g[NV];
    ln_load_row<T, NV>(a, row, lane, ph, thr_pre, ikp, v, mask);
    const float mean = mean_in[row], rstd = rstd_in[row];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * 4;
      g[k] = {0.f, 0.f, 0.f, 0.f};
      if (c < a.cols) {
        f4 d = load4<T>(dy + (long long)row * lddy + c);
        if (a.p_post > 0.f) {
          const f4 m = dropout_scale4(ph, (unsigned long long)row * a.cols + c, a.stream_post, thr_post, ikq);
          d.x *= m.x; d.y *= m.y; d.z *= m.z; d.w *= m.w;
        }
        const f4 gm = load4<T>(a.gamma + c);
        // normalised input
        v[k] = {(v[k].x - mean) * rstd, (v[k].y - mean) * rstd, (v[k].z - mean) * rstd, (v[k].w - mean) * rstd};
        acc_g[k].x += d.x * v[k].x; acc_g[k].y += d.y * v[k].y; acc_g[k].z += d.z * v[k].z; acc_g[k].w += d.w * v[k].w;
        acc_b[k].x += d.x; acc_b[k].y += d.y; acc_b[k].z += d.z; acc_b[k].w += d.w;
        g[k] = {d.x * gm.x, d.y * gm.y, d.z * gm.z, d.w * gm.w};
        s1 += g[k].x + g[k].y + g[k].z + g[k].w;
        s2 += g[k].x * v[k].x + g[k].y * v[k].y + g[k].z * v[k].z + g[k].w * v[k].w;
      }
    }
    const float c1 = warp_sum(s1) * inv_cols, c2 = warp_sum(s2) * inv_cols;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * 4;
      if (c < a.cols) {
        f4 o = {rstd * (g[k].x - c1 - v[k].x * c2), rstd * (g[k].y - c1 - v[k].y * c2),
                rstd * (g[k].z - c1 - v[k].z * c2), rstd * (g[k].w - c1 - v[k].w * c2)};
        if (ds) store4<T>(ds + (long long)row * ldds + c, o);
        o.x *= mask[k].x; o.y *= mask[k].y; o.z *= mask[k].z; o.w *= mask[k].w;   // through the pre-dropout
        if (dz) store4<T>(dz + (long long)row * lddz + c, o);
        acc_bias[k].x += o.x; acc_bias[k].y += o.y; acc_bias[k].z += o.z; acc_bias[k].w += o.w;
      }
    }
  }

  // block fold: three quantities, one after the other through the same smem
  float* out = scratch + (long long)blockIdx.x * 3 * a.cols;
#pragma unroll 1
  for (int qn = 0; qn < 3; ++qn) {
    if (qn == 2 && !want_dbias) break;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * 4;
      if (c < a.cols) {
        const f4 t = qn == 0 ? acc_g[k] : (qn == 1 ? acc_b[k] : acc_bias[k]);
        *reinterpret_cast<float4*>(red + warp * a.cols + c) = make_float4(t.x, t.y, t.z, t.w);
      }
    }
    __syncthreads();
    for (int c = threadIdx.x; c < a.cols; c += kLnThreads) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < kLnWarps; ++w) t += red[w * a.cols + c];
      out[qn * a.cols + c] = t;
    }
  }
}


struct LnDirect {
  float* dgamma; float* dbeta; float* dbias;
};

// Two warps per row: halves the per-thread column accumulators (register pressure was capping the one-warp
// version at one block per SM).  Block = 8 warps = 4 rows per iteration; row statistics are combined through
// shared memory (double-buffered by iteration parity -> one __syncthreads per iteration).  cols % 256 == 0.
// OCC = CTAs per SM the register allocation is held to: 2 -> 128 registers (no spills), 3 -> 80 registers and ~70 bytes
// of spills per thread but half as many warps again to hide the load -> reduce -> store chain behind.  Measured on the
// BERT-base step (B200, 30 steps, twice each): 7.670 ms with 2, 7.712 ms with 3 — the spills cost more than the extra
// warps hide — so 2 is the default and B200_LN_BWD_OCC=3 selects the other instantiation.
// PIPE: as in the forward kernel, x / residual / dy of the NEXT group of four rows are copied to shared memory with
// cp.async (each thread its own vectors) while the current group is reduced and stored.
template <typename T, int NVH, int OCC, bool PIPE>
__global__ void __launch_bounds__(kLnThreads, OCC)
ln_bwd2_kernel(const LnCommon<T> a, const T* __restrict__ dy, long long lddy, const float* __restrict__ mean_in,
               const float* __restrict__ rstd_in, T* __restrict__ ds, long long ldds, T* __restrict__ dz, long long lddz,
               float* __restrict__ scratch, int want_dbias, const LnDirect direct) {
  pdl_trigger();
  pdl_wait();
  extern __shared__ float red[];            // [4 row-slots][cols] for the final fold; PIPE: then R [2][3][NVH][kLnThreads]
  __shared__ float2 part[2][4][2];          // [parity][row slot][half]
  using R = typename Raw4<T>::type;
  R* const buf = reinterpret_cast<R*>(red + 4 * a.cols);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int slot = warp >> 1, half = warp & 1;
  const unsigned long long seed = a.seed_dev ? *a.seed_dev : 0ull;
  const Philox ph(seed);
  const uint32_t thr_pre = dropout_threshold(a.p_pre), thr_post = dropout_threshold(a.p_post);
  const float ikp = a.p_pre > 0.f ? 1.f / (1.f - a.p_pre) : 1.f;
  const float ikq = a.p_post > 0.f ? 1.f / (1.f - a.p_post) : 1.f;
  const float inv_cols = 1.f / (float)a.cols;
  const int cbase = half * (a.cols >> 1);
  f4 acc_g[NVH], acc_b[NVH], acc_bias[NVH];
#pragma unroll
  for (int k = 0; k < NVH; ++k) acc_g[k] = acc_b[k] = acc_bias[k] = {0.f, 0.f, 0.f, 0.f};

  auto issue = [&](int r, int stage) {
#pragma unroll
    for (int k = 0; k < NVH; ++k) {
      const int c = cbase + (k * 32 + lane) * 4;
      cp_async<sizeof(R)>(buf + ((stage * 3 + 0) * NVH + k) * kLnThreads + threadIdx.x, a.x + (long long)r * a.ldx + c);
      if (a.residual)
        cp_async<sizeof(R)>(buf + ((stage * 3 + 1) * NVH + k) * kLnThreads + threadIdx.x, a.residual + (long long)r * a.ldr + c);
      cp_async<sizeof(R)>(buf + ((stage * 3 + 2) * NVH + k) * kLnThreads + threadIdx.x, dy + (long long)r * lddy + c);
    }
  };
  const int step = gridDim.x * 4;
  if constexpr (PIPE) {
    if (blockIdx.x * 4 + slot < a.rows) issue(blockIdx.x * 4 + slot, 0);
    cp_async_commit();
  }
  int it = 0;
  for (int base = blockIdx.x * 4; base < a.rows; base += step, ++it) {
    const int row = base + slot;
    const bool live = row < a.rows;
    f4 v[NVH], g[NVH];
    uint32_t mbits = 0xFFFFFFFFu;            // pre-dropout keep bits, 4 per vector
    float s1 = 0.f, s2 = 0.f, mean = 0.f, rstd = 0.f;
    if constexpr (PIPE) {
      if (row + step < a.rows) issue(row + step, (it + 1) & 1);
      cp_async_commit();                     // one group per iteration, empty or not
      cp_async_wait<1>();
    }
    const int st = it & 1;
    if (live) {
      mean = mean_in[row];
      rstd = rstd_in[row];
#pragma unroll
      for (int k = 0; k < NVH; ++k) {
        const int c = cbase + (k * 32 + lane) * 4;
        f4 t = PIPE ? Raw4<T>::cvt(buf[((st * 3 + 0) * NVH + k) * kLnThreads + threadIdx.x])
                    : load4<T>(a.x + (long long)row * a.ldx + c);
        if (a.bias) {
          const f4 b = load4<T>(a.bias + c);
          t.x += b.x; t.y += b.y; t.z += b.z; t.w += b.w;
        }
        if (a.p_pre > 0.f) {
          const f4 m = dropout_scale4(ph, (unsigned long long)row * a.cols + c, a.stream_pre, thr_pre, ikp);
          t.x *= m.x; t.y *= m.y; t.z *= m.z; t.w *= m.w;
          const uint32_t bits = (m.x != 0.f ? 1u : 0u) | (m.y != 0.f ? 2u : 0u) | (m.z != 0.f ? 4u : 0u) | (m.w != 0.f ? 8u : 0u);
          mbits = (mbits & ~(0xFu << (4 * k))) | (bits << (4 * k));
        }
        if (a.residual) {
          const f4 r = PIPE ? Raw4<T>::cvt(buf[((st * 3 + 1) * NVH + k) * kLnThreads + threadIdx.x])
                            : load4<T>(a.residual + (long long)row * a.ldr + c);
          t.x += r.x; t.y += r.y; t.z += r.z; t.w += r.w;
        }
        f4 d = PIPE ? Raw4<T>::cvt(buf[((st * 3 + 2) * NVH + k) * kLnThreads + threadIdx.x])
                    : load4<T>(dy + (long long)row * lddy + c);
        if (a.p_post > 0.f) {
          const f4 m = dropout_scale4(ph, (unsigned long long)row * a.cols + c, a.stream_post, thr_post, ikq);
          d.x *= m.x; d.y *= m.y; d.z *= m.z; d.w *= m.w;
        }
        const f4 gm = load4<T>(a.gamma + c);
        v[k] = {(t.x - mean) * rstd, (t.y - mean) * rstd, (t.z - mean) * rstd, (t.w - mean) * rstd};
        acc_g[k].x += d.x * v[k].x; acc_g[k].y += d.y * v[k].y; acc_g[k].z += d.z * v[k].z; acc_g[k].w += d.w * v[k].w;
        acc_b[k].x += d.x; acc_b[k].y += d.y; acc_b[k].z += d.z; acc_b[k].w += d.w;
        g[k] = {d.x * gm.x, d.y * gm.y, d.z * gm.z, d.w * gm.w};
        s1 += g[k].x + g[k].y + g[k].z + g[k].w;
        s2 += g[k].x * v[k].x + g[k].y * v[k].y + g[k].z * v[k].z + g[k].w * v[k].w;
      }
    }
    s1 = warp_sum(s1);
    s2 = warp_sum(s2);
    if (lane == 0) part[it & 1][slot][half] = make_float2(s1, s2);
    __syncthreads();
    if (live) {
      const float2 p0 = part[it & 1][slot][0], p1 = part[it & 1][slot][1];
      const float c1 = (p0.x + p1.x) * inv_cols, c2 = (p0.y + p1.y) * inv_cols;
#pragma unroll
      for (int k = 0; k < NVH; ++k) {
        const int c = cbase + (k * 32 + lane) * 4;
        f4 o = {rstd * (g[k].x - c1 - v[k].x * c2), rstd * (g[k].y - c1 - v[k].y * c2),
                rstd * (g[k].z - c1 - v[k].z * c2), rstd * (g[k].w - c1 - v[k].w * c2)};
        if (ds) store4<T>(ds + (long long)row * ldds + c, o);
        if (a.p_pre > 0.f) {
          const uint32_t b = (mbits >> (4 * k)) & 0xFu;
          o.x = (b & 1u) ? o.x * ikp : 0.f; o.y = (b & 2u) ? o.y * ikp : 0.f;
          o.z = (b & 4u) ? o.z * ikp : 0.f; o.w = (b & 8u) ? o.w * ikp : 0.f;
        }
        if (dz) store4<T>(dz + (long long)row * lddz + c, o);
        acc_bias[k].x += o.x; acc_bias[k].y += o.y; acc_bias[k].z += o.z; acc_bias[k].w += o.w;
      }
    }
  }

  // block fold; then either per-block partials to scratch (deterministic two-pass, finalize kernel follows) or,
  // when direct != nullptr, fp32 reductions straight into the parameter gradients (one launch less)
  float* out = scratch ? scratch + (long long)blockIdx.x * 3 * a.cols : nullptr;
#pragma unroll 1
  for (int qn = 0; qn < 3; ++qn) {
    if (qn == 2 && !want_dbias) break;
    __syncthreads();
#pragma unroll
    for (int k = 0; k < NVH; ++k) {
      const int c = cbase + (k * 32 + lane) * 4;
      const f4 t = qn == 0 ? acc_g[k] : (qn == 1 ? acc_b[k] : acc_bias[k]);
      *reinterpret_cast<float4*>(red + slot * a.cols + c) = make_float4(t.x, t.y, t.z, t.w);
    }
    __syncthreads();
    float* dst = qn == 0 ? direct.dgamma : (qn == 1 ? direct.dbeta : direct.dbias);
    for (int c = threadIdx.x; c < a.cols; c += kLnThreads) {
      const float t = red[c] + red[a.cols + c] + red[2 * a.cols + c] + red[3 * a.cols + c];
      if (out) out[qn * a.cols + c] = t;
      else atomicAdd(dst + c, t);
    }
  }
}

// dgamma/dbeta/dbias[c] += sum_b scratch[b][q][c]; one block = 32 columns x 8 row-lanes
__global__ void __launch_bounds__(256)
ln_bwd_finalize(const float* __restrict__ scratch, int nblocks, int cols, float* __restrict__ dgamma,
                float* __restrict__ dbeta, float* __restrict__ dbias) {
  __shared__ float red[8][33];
  const int q = blockIdx.y;
  float* dst = q == 0 ? dgamma : (q == 1 ? dbeta : dbias);
  if (dst == nullptr) return;
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
  const int c = blockIdx.x * 32 + lane;
  float t = 0.f;
  if (c < cols)
    for (int b = w; b < nblocks; b += 8) t += scratch[((long long)b * 3 + q) * cols + c];
  red[w][lane] = t;
  __syncthreads();
  if (w == 0 && c < cols) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[i][lane];
    dst[c] += s;
  }
}

template <typename T>
static LnCommon<T> make_common(const b200_ln_args* a) {
  LnCommon<T> c;
  c.x = (const T*)a->x; c.ldx = a->ldx;
  c.bias = (const T*)a->bias;
  c.residual = (const T*)a->residual; c.ldr = a->ldr;
  c.gamma = (const T*)a->gamma; c.beta = (const T*)a->beta;
  c.eps = a->eps; c.p_pre = a->p_pre; c.p_post = a->p_post;
  c.seed_dev = a->seed_dev; c.stream_pre = a->stream_pre; c.stream_post = a->stream_post;
  c.rows = a->rows; c.cols = a->cols;
  return c;
}

static int check_ln(const b200_ln_args* a) {
  if (!a || !a->x || !a->gamma || !a->beta) return fail(B200_ERR_ARG, "layernorm: null pointer");
  if (a->cols <= 0 || a->cols > 1024 || (a->cols & 3)) return fail(B200_ERR_UNSUPPORTED, "layernorm: cols must be a multiple of 4 and <= 1024 (got %d)", a->cols);
  if ((a->p_pre > 0.f || a->p_post > 0.f) && !a->seed_dev) return fail(B200_ERR_ARG, "layernorm: dropout needs seed_dev");
  const int es = a->dtype == 0 ? 4 : 2;
  if ((a->ldx * es) % (4 * es) || (a->residual && (a->ldr * es) % (4 * es))) return fail(B200_ERR_ARG, "layernorm: pitches must be multiples of 4 elements");
  return 0;
}

// Streaming (cp.async) variants: bit 0 = forward, bit 1 = backward, -1 = automatic; CTAs per SM of each (the grid is that
// many times the SM count, so a warp owns several rows and the prefetch has something to run ahead of).
// B200_LN_PIPE / B200_LN_PIPE_FWD_CTAS / B200_LN_PIPE_BWD_CTAS set the start-up values, b200_ln_config changes them
// at run time.  Measured on B200 (tools/gpu_probe_ln.py, 4096 rows, dropout + bias + residual, us per call):
//   fp32 x 768:   forward 11.8 direct / 13.7 streaming;  backward 19.5 direct / 15.6 streaming (18.7 with 1 CTA/SM)
//   bf16 x 1024:  forward 15.1 / 14.4;                    backward 15.8 / 16.5
// and in the BERT-base TF32 step 7.765 -> 7.659 ms with the streaming backward only (7.82 with both).  The forward
// kernel already keeps 28 warps per SM in flight and gains nothing; the backward kernel is held to 16 warps by its
// registers and gains the memory-level parallelism it lacks.  Automatic = streaming backward for fp32 storage.
struct LnConfig {
  int pipe, fwd_ctas, bwd_ctas;
};
static LnConfig& ln_config() {
  static LnConfig c = [] {
    auto geti = [](const char* n, int d) { const char* e = getenv(n); return e && e[0] ? atoi(e) : d; };
    return LnConfig{geti("B200_LN_PIPE", -1), geti("B200_LN_PIPE_FWD_CTAS", 2), geti("B200_LN_PIPE_BWD_CTAS", 2)};
  }();
  return c;
}

template <typename K>
static int ln_allow_smem(K kernel, size_t bytes) {
  // 40 KB: the kernels also hold a little static shared memory, and the 48 KB default limit covers both
  if (bytes > 40 * 1024) B200_CUDA(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes));
  return 0;
}

template <typename T, int NV>
static int ln_fwd_launch(const b200_ln_args* a, const LnCommon<T>& c, cudaStream_t st) {
  const LnConfig cfg = ln_config();
  int blocks = (a->rows + kLnWarps - 1) / kLnWarps;
  const size_t pipe_smem = (size_t)2 * 2 * NV * kLnThreads * sizeof(typename Raw4<T>::type);
  if (cfg.pipe > 0 && (cfg.pipe & 1) && blocks > num_sms()) {   // fewer rows than warps of one wave: nothing to run ahead of
    const int ctas = cfg.fwd_ctas < 1 ? 1 : (cfg.fwd_ctas > 2 ? 2 : cfg.fwd_ctas);
    if (blocks > num_sms() * ctas) blocks = num_sms() * ctas;
    if (int rc = ln_allow_smem(ln_fwd_kernel<T, NV, true>, pipe_smem)) return rc;
    B200_CUDA(launch_pdl(ln_fwd_kernel<T, NV, true>, dim3(blocks), dim3(kLnThreads), pipe_smem, st, 1, c, (T*)a->y, a->ldy, a->mean, a->rstd));
    B200_LAUNCHED("ln_fwd_kernel<pipe>");
    return 0;
  }
  const int cap = num_sms() * 8;
  if (blocks > cap) blocks = cap;
  B200_CUDA(launch_pdl(ln_fwd_kernel<T, NV, false>, dim3(blocks), dim3(kLnThreads), 0, st, 1, c, (T*)a->y, a->ldy, a->mean, a->rstd));
  B200_LAUNCHED("ln_fwd_kernel");
  return 0;
}

template <typename T>
static int ln_fwd_t(const b200_ln_args* a, cudaStream_t st) {
  const LnCommon<T> c = make_common<T>(a);
  if (a->cols <= 256) return ln_fwd_launch<T, 2>(a, c, st);
  if (a->cols <= 768) return ln_fwd_launch<T, 6>(a, c, st);
  return ln_fwd_launch<T, 8>(a, c, st);
}

static int ln_bwd_blocks(int rows) {
  int blocks = (rows + kLnWarps - 1) / kLnWarps;
  if (blocks > kLnBwdMaxBlocks) blocks = kLnBwdMaxBlocks;
  return blocks < 1 ? 1 : blocks;
}

template <typename T>
static int ln_bwd_t(const b200_ln_args* a, cudaStream_t st) {
  const LnCommon<T> c = make_common<T>(a);
  if ((a->cols & 255) == 0) {
    // two warps per row
    static const int occ = [] { const char* e = getenv("B200_LN_BWD_OCC"); const int v = e ? atoi(e) : 2; return v == 3 ? 3 : 2; }();
    const LnConfig cfg = ln_config();
    int blocks = (a->rows + 3) / 4;
    const bool pipe = (cfg.pipe < 0 ? sizeof(T) == 4 : (cfg.pipe & 2) != 0) && blocks > num_sms();
    int cap = occ == 2 ? kLnBwdMaxBlocks : kLnBwdMaxBlocks3;
    if (pipe) cap = num_sms() * (cfg.bwd_ctas < 1 ? 1 : (cfg.bwd_ctas > 2 ? 2 : cfg.bwd_ctas));
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    const size_t smem2 = (size_t)4 * a->cols * sizeof(float);
    const int want = a->dbias != nullptr;
    const bool direct_mode = a->scratch == nullptr;
    const LnDirect direct{a->dgamma, a->dbeta, a->dbias};
#define LN_BWD2_GO(KERN, SMEM)                                                                                        \
  do {                                                                                                                \
    if (int rc = ln_allow_smem(KERN, SMEM)) return rc;                                                                \
    B200_CUDA(launch_pdl(KERN, dim3(blocks), dim3(kLnThreads), SMEM, st, 1, c, (const T*)a->dy, a->lddy, a->mean,     \
                         a->rstd, (T*)a->ds, a->ldds, (T*)a->dz, a->lddz, (float*)a->scratch, want, direct));         \
  } while (0)
#define LN_BWD2(NVH)                                                                                                  \
  do {                                                                                                                \
    const size_t pipe_smem = smem2 + (size_t)2 * 3 * NVH * kLnThreads * sizeof(typename Raw4<T>::type);               \
    if (pipe) LN_BWD2_GO((ln_bwd2_kernel<T, NVH, 2, true>), pipe_smem);                                               \
    else if (occ == 2) LN_BWD2_GO((ln_bwd2_kernel<T, NVH, 2, false>), smem2);                                         \
    else LN_BWD2_GO((ln_bwd2_kernel<T, NVH, 3, false>), smem2);                                                       \
  } while (0)
    switch (a->cols / 256) {
      case 1: LN_BWD2(1); break;
      case 2: LN_BWD2(2); break;
      case 3: LN_BWD2(3); break;
      default: LN_BWD2(4); break;
    }
#undef LN_BWD2
#undef LN_BWD2_GO
    B200_LAUNCHED(pipe ? "ln_bwd2_kernel<pipe>" : "ln_bwd2_kernel");
    if (direct_mode) return 0;
    dim3 grid2((a->cols + 31) / 32, 3);
    ln_bwd_finalize<<<grid2, 256, 0, st>>>((const float*)a->scratch, blocks, a->cols, a->dgamma, a->dbeta, a->dbias);
    B200_LAUNCHED("ln_bwd_finalize");
    return 0;
  }
  const int blocks = ln_bwd_blocks(a->rows);
  const size_t smem = (size_t)kLnWarps * a->cols * sizeof(float);
  const int want_dbias = a->dbias != nullptr;
#define LN_BWD(NV)                                                                                                   \
  ln_bwd_kernel<T, NV><<<blocks, kLnThreads, smem, st>>>(c, (const T*)a->dy, a->lddy, a->mean, a->rstd, (T*)a->ds,    \
                                                         a->ldds, (T*)a->dz, a->lddz, (float*)a->scratch, want_dbias)
  if (a->cols <= 256) LN_BWD(2);
  else if (a->cols <= 768) LN_BWD(6);
  else LN_BWD(8);
#undef LN_BWD
  B200_LAUNCHED("ln_bwd_kernel");
  dim3 grid((a->cols + 31) / 32, 3);
  ln_bwd_finalize<<<grid, 256, 0, st>>>((const float*)a->scratch, blocks, a->cols, a->dgamma, a->dbeta, a->dbias);
  B200_LAUNCHED("ln_bwd_finalize");
  return 0;
}

}  // namespace b200

using namespace b200;

extern "C" long long b200_ln_bwd_scratch_bytes(int cols) { return (long long)kLnBwdMaxBlocks3 * 3 * cols * 4; }

// Select the streaming (cp.async) LayerNorm kernels at run time: pipe bit 0 = forward, bit 1 = backward; *_ctas = CTAs
// per SM of their grids (1 or 2).  pipe = -1 selects the automatic choice, pipe < -1 / ctas <= 0 leave a field as it
// is.  Not thread-safe against running launches.
extern "C" int b200_ln_config(int pipe, int fwd_ctas, int bwd_ctas) {
  LnConfig& c = ln_config();
  if (pipe >= 0) c.pipe = pipe & 3;
  else if (pipe == -1) c.pipe = -1;
  if (fwd_ctas > 0) c.fwd_ctas = fwd_ctas;
  if (bwd_ctas > 0) c.bwd_ctas = bwd_ctas;
  return c.pipe;
}

extern "C" int b200_ln_fwd(const b200_ln_args* a, void* stream) {
  int rc = check_ln(a);
  if (rc) return rc;
  if (!a->y || !a->mean || !a->rstd) return fail(B200_ERR_ARG, "layernorm fwd: null output");
  return a->dtype == 0 ? ln_fwd_t<float>(a, (cudaStream_t)stream) : ln_fwd_t<__nv_bfloat16>(a, (cudaStream_t)stream);
}

extern "C" int b200_ln_bwd(const b200_ln_args* a, void* stream) {
  int rc = check_ln(a);
  if (rc) return rc;
  if (!a->dy || !a->mean || !a->rstd || !a->dgamma || !a->dbeta)
    return fail(B200_ERR_ARG, "layernorm bwd: null pointer");
  if (!a->scratch && (a->cols & 255))
    return fail(B200_ERR_ARG, "layernorm bwd: scratch may be omitted only when cols is a multiple of 256");
  if (a->scratch && a->scratch_bytes < b200_ln_bwd_scratch_bytes(a->cols)) return fail(B200_ERR_ARG, "layernorm bwd: scratch too small");
  return a->dtype == 0 ? ln_bwd_t<float>(a, (cudaStream_t)stream) : ln_bwd_t<__nv_bfloat16>(a, (cudaStream_t)stream);
}
