// attention_kernels.cu — softmax (+ additive padding mask + dropout) forward / backward over the score
// tensor produced by the batched tcgen05 GEMM (S = QK^T/sqrt(d)), one warp per (batch, head, query) row.
//
// Replaces: scores / sqrt(d) + mask -> nn.Softmax -> nn.Dropout and their autograd duals
// (hetseq/bert_modeling.py:361-371).  The mask is built from the 0/1 attention mask exactly like the
// reference's (1 - mask) * -10000 (bert_modeling.py:817-825).
#include "common.h"
#include "vec_utils.cuh"
#include "../include/b200seq.h"

#include <cfloat>

namespace b200 {

// rows = B*heads*L; row r -> b = r / (heads*L); columns [0,L) valid, [L,Lp) padding (written as 0)
template <typename T, int NV>
__global__ void __launch_bounds__(256)
softmax_fwd_kernel(T* __restrict__ S, T* __restrict__ Pdrop, const long long* __restrict__ attn_mask, int rows, int heads,
                   int L, int Lp, float p_drop, const unsigned long long* __restrict__ seed_dev, unsigned long long stream_id) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const Philox ph(seed_dev ? *seed_dev : 0ull);
  const uint32_t thr = dropout_threshold(p_drop);
  const float ik = p_drop > 0.f ? 1.f / (1.f - p_drop) : 1.f;
  for (int row = warp; row < rows; row += nwarps) {
    const int b = row / (heads * L);
    T* s = S + (long long)row * Lp;
    f4 v[NV];
    float m = -FLT_MAX;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * 4;
      v[k] = {-FLT_MAX, -FLT_MAX, -FLT_MAX, -FLT_MAX};
      if (c < Lp) {
        const f4 t = load4<T>(s + c);
        float* vp = &v[k].x;
        const float* tp = &t.x;
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          if (c + e < L) {
            const float add = attn_mask ? (1.f - (float)attn_mask[(long long)b * L + c + e]) * -10000.f : 0.f;
            vp[e] = tp[e] + add;
            m = fmaxf(m, vp[e]);
          }
        }
      }
    }
    m = warp_max(m);
    float sum = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      float* vp = &v[k].x;
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        vp[e] = vp[e] == -FLT_MAX ? 0.f : __expf(vp[e] - m);
        sum += vp[e];
      }
    }
    const float inv = 1.f / warp_sum(sum);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * 4;
      if (c < Lp) {
        f4 o = {v[k].x * inv, v[k].y * inv, v[k].z * inv, v[k].w * inv};
        store4<T>(s + c, o);
        if (Pdrop) {
          const f4 d = dropout_scale4(ph, (unsigned long long)row * Lp + c, stream_id, thr, ik);
          store4<T>(Pdrop + (long long)row * Lp + c, f4{o.x * d.x, o.y * d.y, o.z * d.z, o.w * d.w});
        }
      }
    }
  }
}

// dS = P * (dPe - sum_k P*dPe),  dPe = dP * dropout_mask / (1-p);  written in place over dP
template <typename T, int NV>
__global__ void __launch_bounds__(256)
softmax_bwd_kernel(const T* __restrict__ P, T* __restrict__ dP, int rows, int L, int Lp, float p_drop,
                   const unsigned long long* __restrict__ seed_dev, unsigned long long stream_id) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  const Philox ph(seed_dev ? *seed_dev : 0ull);
  const uint32_t thr = dropout_threshold(p_drop);
  const float ik = p_drop > 0.f ? 1.f / (1.f - p_drop) : 1.f;
  for (int row = warp; row < rows; row += nwarps) {
    f4 p[NV], g[NV];
    float dot = 0.f;
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * 4;
      p[k] = g[k] = {0.f, 0.f, 0.f, 0.f};
      if (c < Lp) {
        p[k] = load4<T>(P + (long long)row * Lp + c);
        g[k] = load4<T>(dP + (long long)row * Lp + c);
        if (p_drop > 0.f) {
          const f4 d = dropout_scale4(ph, (unsigned long long)row * Lp + c, stream_id, thr, ik);
          g[k] = {g[k].x * d.x, g[k].y * d.y, g[k].z * d.z, g[k].w * d.w};
        }
        dot += p[k].x * g[k].x + p[k].y * g[k].y + p[k].z * g[k].z + p[k].w * g[k].w;
      }
    }
    dot = warp_sum(dot);
#pragma unroll
    for (int k = 0; k < NV; ++k) {
      const int c = (k * 32 + lane) * 4;
      if (c < Lp)
        store4<T>(dP + (long long)row * Lp + c, f4{p[k].x * (g[k].x - dot), p[k].y * (g[k].y - dot),
                                                   p[k].z * (g[k].z - dot), p[k].w * (g[k].w - dot)});
    }
  }
}

static int rows_grid(int rows) {
  int blocks = (rows + 7) / 8;
  const int cap = num_sms() * 16;
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : blocks;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_softmax_fwd(int dtype, void* S, void* Pdrop, const long long* attn_mask, int B, int heads, int L,
                                int Lp, float p_drop, const unsigned long long* seed_dev, unsigned long long stream_id,
                                void* stream) {
  if (!S) return fail(B200_ERR_ARG, "softmax_fwd: null pointer");
  if (Lp > 512 || (Lp & 3) || L > Lp) return fail(B200_ERR_UNSUPPORTED, "softmax_fwd: need L <= Lp <= 512, Lp %% 4 == 0");
  if (p_drop > 0.f && (!seed_dev || !Pdrop)) return fail(B200_ERR_ARG, "softmax_fwd: dropout needs seed_dev and Pdrop");
  const int rows = B * heads * L;
  cudaStream_t st = (cudaStream_t)stream;
#define SM_FWD(T, NV) softmax_fwd_kernel<T, NV><<<rows_grid(rows), 256, 0, st>>>((T*)S, (T*)(p_drop > 0.f ? Pdrop : nullptr), attn_mask, rows, heads, L, Lp, p_drop, seed_dev, stream_id)
  if (dtype == 0) { if (Lp <= 128) SM_FWD(float, 1); else if (Lp <= 256) SM_FWD(float, 2); else SM_FWD(float, 4); }
  else { if (Lp <= 128) SM_FWD(__nv_bfloat16, 1); else if (Lp <= 256) SM_FWD(__nv_bfloat16, 2); else SM_FWD(__nv_bfloat16, 4); }
#undef SM_FWD
  B200_LAUNCHED("softmax_fwd_kernel");
  return 0;
}

extern "C" int b200_softmax_bwd(int dtype, const void* P, void* dP, int B, int heads, int L, int Lp, float p_drop,
                                const unsigned long long* seed_dev, unsigned long long stream_id, void* stream) {
  if (!P || !dP) return fail(B200_ERR_ARG, "softmax_bwd: null pointer");
  if (Lp > 512 || (Lp & 3) || L > Lp) return fail(B200_ERR_UNSUPPORTED, "softmax_bwd: need L <= Lp <= 512, Lp %% 4 == 0");
  if (p_drop > 0.f && !seed_dev) return fail(B200_ERR_ARG, "softmax_bwd: dropout needs seed_dev");
  const int rows = B * heads * L;
  cudaStream_t st = (cudaStream_t)stream;
#define SM_BWD(T, NV) softmax_bwd_kernel<T, NV><<<rows_grid(rows), 256, 0, st>>>((const T*)P, (T*)dP, rows, L, Lp, p_drop, seed_dev, stream_id)
  if (dtype == 0) { if (Lp <= 128) SM_BWD(float, 1); else if (Lp <= 256) SM_BWD(float, 2); else SM_BWD(float, 4); }
  else { if (Lp <= 128) SM_BWD(__nv_bfloat16, 1); else if (Lp <= 256) SM_BWD(__nv_bfloat16, 2); else SM_BWD(__nv_bfloat16, 4); }
#undef SM_BWD
  B200_LAUNCHED("softmax_bwd_kernel");
  return 0;
}
