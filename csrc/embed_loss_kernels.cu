// embed_loss_kernels.cu — embedding gather/scatter, masked-row compaction, softmax cross-entropy and the
// tiny-N classification heads (NSP: N=2, token classification: N=num_labels) for sm_100a.
//
// Reference paths replaced: BertEmbeddings (hetseq/bert_modeling.py:306-320), the full-vocabulary
// CrossEntropyLoss(ignore_index=-1) over every position (:899-902), the NSP head + CE (:580,902) and the
// token-classification head with boolean-mask gather + CE (:1221-1245, which costs a host sync).
#include "common.h"
#include "vec_utils.cuh"
#include "../include/b200seq.h"

#include <cfloat>
#include <mutex>

namespace b200 {

__device__ __forceinline__ void red_add_f4(float* addr, const f4& v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w)
               : "memory");
}

// ---- embeddings --------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256)
embed_sum_kernel(const long long* __restrict__ ids, const long long* __restrict__ tts, const T* __restrict__ W,
                 const T* __restrict__ P, const T* __restrict__ TT, T* __restrict__ out, int rows, int L, int H,
                 int vocab, int ntypes) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int row = warp; row < rows; row += nwarps) {
    long long id = ids[row];
    long long tt = tts ? tts[row] : 0;
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    tt = tt < 0 ? 0 : (tt >= ntypes ? ntypes - 1 : tt);
    const int pos = row % L;
    for (int c = lane * 4; c < H; c += 128) {
      const f4 a = load4<T>(W + id * H + c), b = load4<T>(P + (long long)pos * H + c), t = load4<T>(TT + tt * H + c);
      store4<T>(out + (long long)row * H + c, f4{a.x + b.x + t.x, a.y + b.y + t.y, a.z + b.z + t.z, a.w + b.w + t.w});
    }
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
embed_scatter_kernel(const long long* __restrict__ ids, const long long* __restrict__ tts, const T* __restrict__ ds,
                     float* __restrict__ gW, float* __restrict__ gP, float* __restrict__ gT, int rows, int L, int H,
                     int vocab, int ntypes) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int row = warp; row < rows; row += nwarps) {
    long long id = ids[row];
    long long tt = tts ? tts[row] : 0;
    id = id < 0 ? 0 : (id >= vocab ? vocab - 1 : id);
    tt = tt < 0 ? 0 : (tt >= ntypes ? ntypes - 1 : tt);
    const int pos = row % L;
    for (int c = lane * 4; c < H; c += 128) {
      const f4 v = load4<T>(ds + (long long)row * H + c);
      red_add_f4(gW + id * H + c, v);
      red_add_f4(gP + (long long)pos * H + c, v);
      red_add_f4(gT + tt * H + c, v);
    }
  }
}

// ---- masked-row compaction (single block, order preserving) -------------------------------------------
// idx_out[j] = j-th row r with labels[r] != ignore (and mask[r] != 0), label_out[j] its label; the tail of
// both arrays is filled with -1 / ignore.  count_out[0] = min(n_valid, cap), count_out[1] = n_valid.
__global__ void __launch_bounds__(1024)
compact_rows_kernel(const long long* __restrict__ labels, const long long* __restrict__ mask, int rows, int ignore,
                    int cap, int* __restrict__ idx_out, long long* __restrict__ label_out, int* __restrict__ count_out) {
  __shared__ int warp_tot[32];
  __shared__ int base;
  if (threadIdx.x == 0) base = 0;
  __syncthreads();
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  for (int start = 0; start < rows; start += 1024) {
    const int r = start + threadIdx.x;
    int valid = 0;
    long long lab = ignore;
    if (r < rows) {
      lab = labels[r];
      valid = (lab != ignore) && (!mask || mask[r] != 0);
    }
    const unsigned bal = __ballot_sync(0xffffffffu, valid);
    const int within = __popc(bal & ((1u << lane) - 1));
    if (lane == 0) warp_tot[warp] = __popc(bal);
    __syncthreads();
    int off = base;
    for (int w = 0; w < warp; ++w) off += warp_tot[w];
    if (valid) {
      const int j = off + within;
      if (j < cap) {
        idx_out[j] = r;
        label_out[j] = lab;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < 32; ++w) t += warp_tot[w];
      base += t;
    }
    __syncthreads();
  }
  const int n = base;
  for (int j = (n < cap ? n : cap) + threadIdx.x; j < cap; j += 1024) {
    idx_out[j] = -1;
    label_out[j] = ignore;
  }
  if (threadIdx.x == 0) {
    count_out[0] = n < cap ? n : cap;
    count_out[1] = n;
  }
}

template <typename T>
__global__ void __launch_bounds__(256)
gather_rows_kernel(const T* __restrict__ x, long long ldx, const int* __restrict__ idx, int n, int H, T* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int i = warp; i < n; i += nwarps) {
    const int r = idx[i];
    for (int c = lane * 4; c < H; c += 128) {
      const f4 v = r >= 0 ? load4<T>(x + (long long)r * ldx + c) : f4{0.f, 0.f, 0.f, 0.f};
      store4<T>(out + (long long)i * H + c, v);
    }
  }
}

// dx[idx[i]] (+)= dy[i]; indices are unique so no atomics are needed.
template <typename T>
__global__ void __launch_bounds__(256)
scatter_rows_kernel(const T* __restrict__ dy, const int* __restrict__ idx, int n, int H, T* __restrict__ dx, long long lddx,
                    int accumulate) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int i = warp; i < n; i += nwarps) {
    const int r = idx[i];
    if (r < 0) continue;
    for (int c = lane * 4; c < H; c += 128) {
      f4 v = load4<T>(dy + (long long)i * H + c);
      if (accumulate) {
        const f4 o = load4<T>(dx + (long long)r * lddx + c);
        v = {v.x + o.x, v.y + o.y, v.z + o.z, v.w + o.w};
      }
      store4<T>(dx + (long long)r * lddx + c, v);
    }
  }
}

// ---- softmax cross-entropy over a wide vocabulary -------------------------------------------------------
// one block per row; online (max, sum) in fp32; labels == ignore produce zero loss / zero gradient
template <typename T>
__global__ void __launch_bounds__(256)
ce_fwd_kernel(const T* __restrict__ logits, long long ld, const long long* __restrict__ labels, int V, int ignore,
              float* __restrict__ lse_out, float* __restrict__ loss_out) {
  __shared__ float sm[8], ss[8];
  const int row = blockIdx.x;
  const long long lab = labels[row];
  if (lab == ignore || lab < 0 || lab >= V) {
    if (threadIdx.x == 0) {
      lse_out[row] = 0.f;
      loss_out[row] = 0.f;
    }
    return;
  }
  const T* p = logits + (long long)row * ld;
  float m = -FLT_MAX, s = 0.f;
  const int V4 = V & ~3;
  for (int c = threadIdx.x * 4; c < V4; c += 1024) {
    const f4 v = load4<T>(p + c);
    const float mx = fmaxf(fmaxf(v.x, v.y), fmaxf(v.z, v.w));
    if (mx > m) {
      s *= __expf(m - mx);
      m = mx;
    }
    s += __expf(v.x - m) + __expf(v.y - m) + __expf(v.z - m) + __expf(v.w - m);
  }
  if (threadIdx.x < V - V4) {
    const float v = load1<T>(p + V4 + threadIdx.x);
    if (v > m) {
      s *= __expf(m - v);
      m = v;
    }
    s += __expf(v - m);
  }
  // combine (m, s) pairs across the block
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float m2 = __shfl_xor_sync(0xffffffffu, m, o), s2 = __shfl_xor_sync(0xffffffffu, s, o);
    const float mm = fmaxf(m, m2);
    s = s * __expf(m - mm) + s2 * __expf(m2 - mm);
    m = mm;
  }
  if ((threadIdx.x & 31) == 0) {
    sm[threadIdx.x >> 5] = m;
    ss[threadIdx.x >> 5] = s;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float M = sm[0], S = ss[0];
    for (int w = 1; w < 8; ++w) {
      const float mm = fmaxf(M, sm[w]);
      S = S * __expf(M - mm) + ss[w] * __expf(sm[w] - mm);
      M = mm;
    }
    const float lse = M + logf(S);
    lse_out[row] = lse;
    loss_out[row] = lse - load1<T>(p + lab);
  }
}

// dlogits = (softmax - onehot) * gout * inv_count  (written in place over the logits)
template <typename T>
__global__ void __launch_bounds__(256)
ce_bwd_kernel(T* __restrict__ logits, long long ld, const long long* __restrict__ labels, int V, int ignore,
              const float* __restrict__ lse, const float* __restrict__ gout, const float* __restrict__ inv_count) {
  const int row = blockIdx.x;
  const long long lab = labels[row];
  T* p = logits + (long long)row * ld;
  const bool dead = (lab == ignore || lab < 0 || lab >= V);
  const float scale = dead ? 0.f : (gout ? *gout : 1.f) * (inv_count ? *inv_count : 1.f);
  const float l = lse[row];
  const int V4 = V & ~3;
  for (int c = threadIdx.x * 4; c < V4; c += 1024) {
    f4 v = {0.f, 0.f, 0.f, 0.f};
    if (!dead) {
      v = load4<T>(p + c);
      v = {__expf(v.x - l) * scale, __expf(v.y - l) * scale, __expf(v.z - l) * scale, __expf(v.w - l) * scale};
      const int d = (int)lab - c;
      if (d == 0) v.x -= scale; else if (d == 1) v.y -= scale; else if (d == 2) v.z -= scale; else if (d == 3) v.w -= scale;
    }
    store4<T>(p + c, v);
  }
  if (threadIdx.x < V - V4) {
    const int c = V4 + threadIdx.x;
    float v = 0.f;
    if (!dead) v = __expf(load1<T>(p + c) - l) * scale - (c == lab ? scale : 0.f);
    store1<T>(p + c, v);
  }
}

// loss[0] = sum(row_loss) / max(count,1); inv_count[0] = 1 / max(count,1); single block, fixed order
__global__ void __launch_bounds__(1024)
ce_finalize_kernel(const float* __restrict__ row_loss, const long long* __restrict__ labels, int rows, int ignore,
                   float* __restrict__ loss, float* __restrict__ inv_count, float* __restrict__ count_out) {
  __shared__ float s_loss[32];
  __shared__ int s_cnt[32];
  float t = 0.f;
  int n = 0;
  for (int r = threadIdx.x; r < rows; r += 1024) {
    t += row_loss[r];
    n += (labels[r] != ignore) ? 1 : 0;
  }
  t = warp_sum(t);
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) n += __shfl_xor_sync(0xffffffffu, n, o);
  if ((threadIdx.x & 31) == 0) {
    s_loss[threadIdx.x >> 5] = t;
    s_cnt[threadIdx.x >> 5] = n;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    float T = 0.f;
    int N = 0;
    for (int w = 0; w < 32; ++w) {
      T += s_loss[w];
      N += s_cnt[w];
    }
    const float inv = 1.f / (float)(N > 0 ? N : 1);
    loss[0] = T * inv;
    if (inv_count) inv_count[0] = inv;
    if (count_out) count_out[0] = (float)N;
  }
}

// ---- tiny-N linear heads (N <= 32): logits fp32 ---------------------------------------------------------
constexpr int kMaxHeadN = 32;

template <typename T>
__global__ void __launch_bounds__(256)
head_fwd_kernel(const T* __restrict__ x, long long ldx, const T* __restrict__ w, const T* __restrict__ b, int rows, int K,
                int N, float* __restrict__ out) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int row = warp; row < rows; row += nwarps) {
    float acc[kMaxHeadN];
#pragma unroll
    for (int n = 0; n < kMaxHeadN; ++n) acc[n] = 0.f;
    for (int c = lane * 4; c < K; c += 128) {
      const f4 v = load4<T>(x + (long long)row * ldx + c);
#pragma unroll
      for (int n = 0; n < kMaxHeadN; ++n) {
        if (n < N) {
          const f4 ww = load4<T>(w + (long long)n * K + c);
          acc[n] += v.x * ww.x + v.y * ww.y + v.z * ww.z + v.w * ww.w;
        }
      }
    }
#pragma unroll
    for (int n = 0; n < kMaxHeadN; ++n) {
      if (n < N) {
        const float t = warp_sum(acc[n]);
        if (lane == 0) out[(long long)row * N + n] = t + (b ? load1<T>(b + n) : 0.f);
      }
    }
  }
}

// small-N softmax CE on fp32 logits: one thread per row; optional active-mask gating (token classification)
__global__ void __launch_bounds__(256)
head_ce_kernel(const float* __restrict__ logits, const long long* __restrict__ labels, const long long* __restrict__ active,
               int rows, int N, int ignore, float* __restrict__ row_loss, long long* __restrict__ eff_labels) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  long long lab = labels[r];
  if (active && active[r] != 1) lab = ignore;
  if (lab < 0 || lab >= N) lab = ignore;
  eff_labels[r] = lab;
  if (lab == ignore) {
    row_loss[r] = 0.f;
    return;
  }
  const float* p = logits + (long long)r * N;
  float m = -FLT_MAX;
  for (int n = 0; n < N; ++n) m = fmaxf(m, p[n]);
  float s = 0.f;
  for (int n = 0; n < N; ++n) s += __expf(p[n] - m);
  row_loss[r] = m + logf(s) - p[lab];
}

// dlogits[r,n] = (softmax - onehot) * gout * inv_count   (0 for ignored rows)
__global__ void __launch_bounds__(256)
head_ce_bwd_kernel(const float* __restrict__ logits, const long long* __restrict__ eff_labels, int rows, int N, int ignore,
                   const float* __restrict__ gout, const float* __restrict__ inv_count, float* __restrict__ dlogits) {
  const int r = blockIdx.x * blockDim.x + threadIdx.x;
  if (r >= rows) return;
  const long long lab = eff_labels[r];
  float* d = dlogits + (long long)r * N;
  if (lab == ignore) {
    for (int n = 0; n < N; ++n) d[n] = 0.f;
    return;
  }
  const float scale = (gout ? *gout : 1.f) * (inv_count ? *inv_count : 1.f);
  const float* p = logits + (long long)r * N;
  float m = -FLT_MAX;
  for (int n = 0; n < N; ++n) m = fmaxf(m, p[n]);
  float s = 0.f;
  for (int n = 0; n < N; ++n) s += __expf(p[n] - m);
  const float inv = 1.f / s;
  for (int n = 0; n < N; ++n) d[n] = (__expf(p[n] - m) * inv - (n == lab ? 1.f : 0.f)) * scale;
}

// dx[r,:] (+)= sum_n dl[r,n] w[n,:]
template <typename T>
__global__ void __launch_bounds__(256)
head_bwd_dx_kernel(const float* __restrict__ dl, const T* __restrict__ w, int rows, int K, int N, T* __restrict__ dx,
                   long long lddx, int accumulate) {
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  const int nwarps = (gridDim.x * blockDim.x) >> 5;
  for (int row = warp; row < rows; row += nwarps) {
    float d[kMaxHeadN];
#pragma unroll
    for (int n = 0; n < kMaxHeadN; ++n) d[n] = n < N ? dl[(long long)row * N + n] : 0.f;
    for (int c = lane * 4; c < K; c += 128) {
      f4 acc = accumulate ? load4<T>(dx + (long long)row * lddx + c) : f4{0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int n = 0; n < kMaxHeadN; ++n) {
        if (n < N) {
          const f4 ww = load4<T>(w + (long long)n * K + c);
          acc.x += d[n] * ww.x; acc.y += d[n] * ww.y; acc.z += d[n] * ww.z; acc.w += d[n] * ww.w;
        }
      }
      store4<T>(dx + (long long)row * lddx + c, acc);
    }
  }
}

// dw[n,:] += sum_r dl[r,n] x[r,:],  db[n] += sum_r dl[r,n].  Block = 64 rows x all K (thread owns 4 columns).
template <typename T>
__global__ void __launch_bounds__(256)
head_bwd_dw_kernel(const float* __restrict__ dl, const T* __restrict__ x, long long ldx, int rows, int K, int N,
                   float* __restrict__ dw, float* __restrict__ db) {
  __shared__ float sdl[64 * kMaxHeadN];
  const int r0 = blockIdx.x * 64;
  const int nr = min(64, rows - r0);
  for (int i = threadIdx.x; i < nr * N; i += blockDim.x) sdl[i] = dl[(long long)r0 * N + i];
  __syncthreads();
  for (int c = threadIdx.x * 4; c < K; c += blockDim.x * 4) {
    f4 acc[kMaxHeadN];
#pragma unroll
    for (int n = 0; n < kMaxHeadN; ++n) acc[n] = {0.f, 0.f, 0.f, 0.f};
    for (int r = 0; r < nr; ++r) {
      const f4 v = load4<T>(x + (long long)(r0 + r) * ldx + c);
#pragma unroll
      for (int n = 0; n < kMaxHeadN; ++n) {
        if (n < N) {
          const float d = sdl[r * N + n];
          acc[n].x += d * v.x; acc[n].y += d * v.y; acc[n].z += d * v.z; acc[n].w += d * v.w;
        }
      }
    }
#pragma unroll
    for (int n = 0; n < kMaxHeadN; ++n)
      if (n < N) red_add_f4(dw + (long long)n * K + c, acc[n]);
  }
  if (db && threadIdx.x < N) {
    float t = 0.f;
    for (int r = 0; r < nr; ++r) t += sdl[r * N + threadIdx.x];
    atomicAdd(db + threadIdx.x, t);
  }
}

// column sums: out[c] += sum_r x[r,c]   (bias gradients of the big GEMMs); two-stage through scratch.
// Block (bx, by): 128 columns [bx*128, +128) x row range by; 8 warps interleave the rows (each warp reads a
// coalesced 512 B row segment), partials meet in shared memory, scratch[by][c] gets the block's sum.
template <typename T>
__global__ void __launch_bounds__(256)
colsum_partial_kernel(const T* __restrict__ x, long long ldx, int rows, int cols, int rows_per_block, float* __restrict__ scratch) {
  __shared__ float4 red[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 128 + lane * 4;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(rows, r0 + rows_per_block);
  f4 acc = {0.f, 0.f, 0.f, 0.f};
  if (c < cols) {
    int r = r0 + warp;
    for (; r + 24 < r1; r += 32) {          // 4 independent loads in flight
      const f4 v0 = load4<T>(x + (long long)r * ldx + c), v1 = load4<T>(x + (long long)(r + 8) * ldx + c);
      const f4 v2 = load4<T>(x + (long long)(r + 16) * ldx + c), v3 = load4<T>(x + (long long)(r + 24) * ldx + c);
      acc.x += (v0.x + v1.x) + (v2.x + v3.x); acc.y += (v0.y + v1.y) + (v2.y + v3.y);
      acc.z += (v0.z + v1.z) + (v2.z + v3.z); acc.w += (v0.w + v1.w) + (v2.w + v3.w);
    }
    for (; r < r1; r += 8) {
      const f4 v = load4<T>(x + (long long)r * ldx + c);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  red[warp][lane] = make_float4(acc.x, acc.y, acc.z, acc.w);
  __syncthreads();
  if (warp == 0 && c < cols) {
    float4 t = red[0][lane];
#pragma unroll
    for (int w = 1; w < 8; ++w) { t.x += red[w][lane].x; t.y += red[w][lane].y; t.z += red[w][lane].z; t.w += red[w][lane].w; }
    *reinterpret_cast<float4*>(scratch + (long long)blockIdx.y * cols + c) = t;
  }
}
// single-pass variant: the block's 128 column sums go straight into out[] with vector reductions at L2
template <typename T>
__global__ void __launch_bounds__(256)
colsum_atomic_kernel(const T* __restrict__ x, long long ldx, int rows, int cols, int rows_per_block, float* __restrict__ out) {
  pdl_trigger();
  pdl_wait();
  __shared__ float4 red[8][32];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int c = blockIdx.x * 128 + lane * 4;
  const int r0 = blockIdx.y * rows_per_block;
  const int r1 = min(rows, r0 + rows_per_block);
  f4 acc = {0.f, 0.f, 0.f, 0.f};
  if (c < cols) {
    int r = r0 + warp;
    for (; r + 24 < r1; r += 32) {
      const f4 v0 = load4<T>(x + (long long)r * ldx + c), v1 = load4<T>(x + (long long)(r + 8) * ldx + c);
      const f4 v2 = load4<T>(x + (long long)(r + 16) * ldx + c), v3 = load4<T>(x + (long long)(r + 24) * ldx + c);
      acc.x += (v0.x + v1.x) + (v2.x + v3.x); acc.y += (v0.y + v1.y) + (v2.y + v3.y);
      acc.z += (v0.z + v1.z) + (v2.z + v3.z); acc.w += (v0.w + v1.w) + (v2.w + v3.w);
    }
    for (; r < r1; r += 8) {
      const f4 v = load4<T>(x + (long long)r * ldx + c);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  red[warp][lane] = make_float4(acc.x, acc.y, acc.z, acc.w);
  __syncthreads();
  if (warp == 0 && c < cols) {
    float4 t = red[0][lane];
#pragma unroll
    for (int w = 1; w < 8; ++w) { t.x += red[w][lane].x; t.y += red[w][lane].y; t.z += red[w][lane].z; t.w += red[w][lane].w; }
    red_add_f4(out + c, f4{t.x, t.y, t.z, t.w});
  }
}

__global__ void __launch_bounds__(256)
colsum_final_kernel(const float* __restrict__ scratch, int nparts, int cols, float* __restrict__ out) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= cols) return;
  float t = 0.f;
  for (int p = 0; p < nparts; ++p) t += scratch[(long long)p * cols + c];
  out[c] += t;
}

// y = x * keep/(1-p), mask regenerated from (seed, stream id, element index); n % 4 == 0
template <typename T>
__global__ void __launch_bounds__(256)
dropout_kernel(const T* __restrict__ x, T* __restrict__ y, long long n, float p, const unsigned long long* __restrict__ seed_dev,
               unsigned long long stream_id) {
  const Philox ph(seed_dev ? *seed_dev : 0ull);
  const uint32_t thr = dropout_threshold(p);
  const float ik = p > 0.f ? 1.f / (1.f - p) : 1.f;
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const f4 v = load4<T>(x + (i << 2));
    const f4 d = dropout_scale4(ph, (unsigned long long)(i << 2), stream_id, thr, ik);
    store4<T>(y + (i << 2), f4{v.x * d.x, v.y * d.y, v.z * d.z, v.w * d.w});
  }
}

// out = dy * act'(aux): mode 1 -> gelu'(aux = pre-activation), mode 2 -> tanh' = 1 - aux^2 (aux = tanh output)
template <typename T>
__global__ void __launch_bounds__(256)
act_bwd_kernel(const T* __restrict__ dy, const T* __restrict__ aux, T* __restrict__ out, long long n, int mode) {
  const long long n4 = n >> 2;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x) {
    const f4 g = load4<T>(dy + (i << 2)), a = load4<T>(aux + (i << 2));
    const float* gp = &g.x; const float* ap = &a.x;
    f4 o; float* op = &o.x;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float x = ap[e];
      float d;
      if (mode == 1) {
        const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752f));
        d = cdf + x * 0.3989422804014327f * __expf(-0.5f * x * x);
      } else {
        d = 1.0f - x * x;
      }
      op[e] = gp[e] * d;
    }
    store4<T>(out + (i << 2), o);
  }
}

static int warp_rows_grid(int rows) {
  int blocks = (rows + 7) / 8;
  const int cap = num_sms() * 8;
  if (blocks > cap) blocks = cap;
  return blocks < 1 ? 1 : blocks;
}

}  // namespace b200

using namespace b200;

#define DISPATCH_T(dtype, CALL)                         \
  do {                                                  \
    if ((dtype) == 0) { using T = float; CALL; }        \
    else { using T = __nv_bfloat16; CALL; }             \
  } while (0)

extern "C" int b200_embed_sum(int dtype, const long long* ids, const long long* tts, const void* W, const void* P,
                              const void* TT, void* out, int rows, int L, int H, int vocab, int ntypes, void* stream) {
  if (!ids || !W || !P || !TT || !out) return fail(B200_ERR_ARG, "embed_sum: null pointer");
  if (H & 3) return fail(B200_ERR_UNSUPPORTED, "embed_sum: hidden size must be a multiple of 4");
  DISPATCH_T(dtype, (embed_sum_kernel<T><<<warp_rows_grid(rows), 256, 0, (cudaStream_t)stream>>>(
                        ids, tts, (const T*)W, (const T*)P, (const T*)TT, (T*)out, rows, L, H, vocab, ntypes)));
  B200_LAUNCHED("embed_sum_kernel");
  return 0;
}

extern "C" int b200_embed_scatter(int dtype, const long long* ids, const long long* tts, const void* ds, float* gW,
                                  float* gP, float* gT, int rows, int L, int H, int vocab, int ntypes, void* stream) {
  if (!ids || !ds || !gW || !gP || !gT) return fail(B200_ERR_ARG, "embed_scatter: null pointer");
  if (H & 3) return fail(B200_ERR_UNSUPPORTED, "embed_scatter: hidden size must be a multiple of 4");
  DISPATCH_T(dtype, (embed_scatter_kernel<T><<<warp_rows_grid(rows), 256, 0, (cudaStream_t)stream>>>(
                        ids, tts, (const T*)ds, gW, gP, gT, rows, L, H, vocab, ntypes)));
  B200_LAUNCHED("embed_scatter_kernel");
  return 0;
}

extern "C" int b200_compact_rows(const long long* labels, const long long* mask, int rows, int ignore, int cap,
                                 int* idx_out, long long* label_out, int* count_out, void* stream) {
  if (!labels || !idx_out || !label_out || !count_out) return fail(B200_ERR_ARG, "compact_rows: null pointer");
  compact_rows_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(labels, mask, rows, ignore, cap, idx_out, label_out, count_out);
  B200_LAUNCHED("compact_rows_kernel");
  return 0;
}

extern "C" int b200_gather_rows(int dtype, const void* x, long long ldx, const int* idx, int n, int H, void* out, void* stream) {
  if (!x || !idx || !out) return fail(B200_ERR_ARG, "gather_rows: null pointer");
  if (H & 3) return fail(B200_ERR_UNSUPPORTED, "gather_rows: width must be a multiple of 4");
  DISPATCH_T(dtype, (gather_rows_kernel<T><<<warp_rows_grid(n), 256, 0, (cudaStream_t)stream>>>((const T*)x, ldx, idx, n, H, (T*)out)));
  B200_LAUNCHED("gather_rows_kernel");
  return 0;
}

extern "C" int b200_scatter_rows(int dtype, const void* dy, const int* idx, int n, int H, void* dx, long long lddx,
                                 int accumulate, void* stream) {
  if (!dy || !idx || !dx) return fail(B200_ERR_ARG, "scatter_rows: null pointer");
  if (H & 3) return fail(B200_ERR_UNSUPPORTED, "scatter_rows: width must be a multiple of 4");
  DISPATCH_T(dtype, (scatter_rows_kernel<T><<<warp_rows_grid(n), 256, 0, (cudaStream_t)stream>>>((const T*)dy, idx, n, H, (T*)dx, lddx, accumulate)));
  B200_LAUNCHED("scatter_rows_kernel");
  return 0;
}

extern "C" int b200_ce_fwd(int dtype, const void* logits, long long ld, const long long* labels, int rows, int V,
                           int ignore, float* lse, float* row_loss, void* stream) {
  if (!logits || !labels || !lse || !row_loss) return fail(B200_ERR_ARG, "ce_fwd: null pointer");
  if (ld & 3) return fail(B200_ERR_ARG, "ce_fwd: logits pitch must be a multiple of 4 elements");
  if (rows <= 0) return 0;
  DISPATCH_T(dtype, (ce_fwd_kernel<T><<<rows, 256, 0, (cudaStream_t)stream>>>((const T*)logits, ld, labels, V, ignore, lse, row_loss)));
  B200_LAUNCHED("ce_fwd_kernel");
  return 0;
}

extern "C" int b200_ce_bwd(int dtype, void* logits, long long ld, const long long* labels, int rows, int V, int ignore,
                           const float* lse, const float* gout, const float* inv_count, void* stream) {
  if (!logits || !labels || !lse) return fail(B200_ERR_ARG, "ce_bwd: null pointer");
  if (rows <= 0) return 0;
  DISPATCH_T(dtype, (ce_bwd_kernel<T><<<rows, 256, 0, (cudaStream_t)stream>>>((T*)logits, ld, labels, V, ignore, lse, gout, inv_count)));
  B200_LAUNCHED("ce_bwd_kernel");
  return 0;
}

extern "C" int b200_ce_finalize(const float* row_loss, const long long* labels, int rows, int ignore, float* loss,
                                float* inv_count, float* count_out, void* stream) {
  if (!row_loss || !labels || !loss) return fail(B200_ERR_ARG, "ce_finalize: null pointer");
  ce_finalize_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(row_loss, labels, rows, ignore, loss, inv_count, count_out);
  B200_LAUNCHED("ce_finalize_kernel");
  return 0;
}

extern "C" int b200_head_fwd(int dtype, const void* x, long long ldx, const void* w, const void* b, int rows, int K, int N,
                             float* out, void* stream) {
  if (!x || !w || !out) return fail(B200_ERR_ARG, "head_fwd: null pointer");
  if (N > kMaxHeadN || N < 1 || (K & 3)) return fail(B200_ERR_UNSUPPORTED, "head_fwd: need 1 <= N <= 32 and K %% 4 == 0");
  DISPATCH_T(dtype, (head_fwd_kernel<T><<<warp_rows_grid(rows), 256, 0, (cudaStream_t)stream>>>((const T*)x, ldx, (const T*)w, (const T*)b, rows, K, N, out)));
  B200_LAUNCHED("head_fwd_kernel");
  return 0;
}

extern "C" int b200_head_ce(const float* logits, const long long* labels, const long long* active, int rows, int N,
                            int ignore, float* row_loss, long long* eff_labels, void* stream) {
  if (!logits || !labels || !row_loss || !eff_labels) return fail(B200_ERR_ARG, "head_ce: null pointer");
  head_ce_kernel<<<(rows + 255) / 256, 256, 0, (cudaStream_t)stream>>>(logits, labels, active, rows, N, ignore, row_loss, eff_labels);
  B200_LAUNCHED("head_ce_kernel");
  return 0;
}

extern "C" int b200_head_ce_bwd(const float* logits, const long long* eff_labels, int rows, int N, int ignore,
                                const float* gout, const float* inv_count, float* dlogits, void* stream) {
  if (!logits || !eff_labels || !dlogits) return fail(B200_ERR_ARG, "head_ce_bwd: null pointer");
  head_ce_bwd_kernel<<<(rows + 255) / 256, 256, 0, (cudaStream_t)stream>>>(logits, eff_labels, rows, N, ignore, gout, inv_count, dlogits);
  B200_LAUNCHED("head_ce_bwd_kernel");
  return 0;
}

extern "C" int b200_head_bwd(int dtype, const float* dl, const void* x, long long ldx, const void* w, int rows, int K, int N,
                             void* dx, long long lddx, int accumulate_dx, float* dw, float* db, void* stream) {
  if (!dl || !x || !w) return fail(B200_ERR_ARG, "head_bwd: null pointer");
  if (N > kMaxHeadN || N < 1 || (K & 3)) return fail(B200_ERR_UNSUPPORTED, "head_bwd: need 1 <= N <= 32 and K %% 4 == 0");
  if (dx) {
    DISPATCH_T(dtype, (head_bwd_dx_kernel<T><<<warp_rows_grid(rows), 256, 0, (cudaStream_t)stream>>>(dl, (const T*)w, rows, K, N, (T*)dx, lddx, accumulate_dx)));
    B200_LAUNCHED("head_bwd_dx_kernel");
  }
  if (dw) {
    DISPATCH_T(dtype, (head_bwd_dw_kernel<T><<<(rows + 63) / 64, 256, 0, (cudaStream_t)stream>>>(dl, (const T*)x, ldx, rows, K, N, dw, db)));
    B200_LAUNCHED("head_bwd_dw_kernel");
  }
  return 0;
}

extern "C" long long b200_colsum_scratch_bytes(int rows, int cols) {
  const int parts = rows >= 2048 ? 64 : (rows >= 256 ? 16 : 1);
  return (long long)parts * cols * 4;
}

// vector loads in the column-sum kernels fetch 4 elements at once: base and pitch must keep every row 16-byte
// (fp32) / 8-byte (bf16) aligned and the pitch must cover a row
static int colsum_check_layout(const char* who, int dtype, const void* x, long long ldx, int cols) {
  const uintptr_t align = dtype == 0 ? 16 : 8;
  if ((reinterpret_cast<uintptr_t>(x) & (align - 1)) != 0) return fail(B200_ERR_ARG, "%s: x must be %d-byte aligned", who, (int)align);
  if (ldx < cols || (ldx & 3)) return fail(B200_ERR_ARG, "%s: row pitch %lld must be a multiple of 4 elements and >= cols (%d)", who, ldx, cols);
  return 0;
}

extern "C" int b200_colsum(int dtype, const void* x, long long ldx, int rows, int cols, float* out, void* scratch,
                           long long scratch_bytes, void* stream) {
  if (!x || !out) return fail(B200_ERR_ARG, "colsum: null pointer");
  if (cols & 3) return fail(B200_ERR_UNSUPPORTED, "colsum: cols must be a multiple of 4");
  if (const int lrc = colsum_check_layout("colsum", dtype, x, ldx, cols)) return lrc;
  const int parts = rows >= 2048 ? 64 : (rows >= 256 ? 16 : 1);
  if (scratch == nullptr) {
    // no scratch given: one kernel, partial sums added to out[] with fp32 reductions (summation order varies)
    if ((reinterpret_cast<uintptr_t>(out) & 15) != 0) return fail(B200_ERR_ARG, "colsum: out must be 16-byte aligned");
    const int rpb0 = (rows + parts - 1) / parts;
    dim3 grid0((cols + 127) / 128, parts);
    DISPATCH_T(dtype, (launch_pdl(colsum_atomic_kernel<T>, grid0, dim3(256), 0, (cudaStream_t)stream, 1, (const T*)x, ldx, rows, cols, rpb0, out)));
    B200_LAUNCHED("colsum_atomic_kernel");
    return 0;
  }
  if (scratch_bytes < (long long)parts * cols * 4) return fail(B200_ERR_ARG, "colsum: scratch too small");
  const int rpb = (rows + parts - 1) / parts;
  dim3 grid((cols + 127) / 128, parts);
  DISPATCH_T(dtype, (colsum_partial_kernel<T><<<grid, 256, 0, (cudaStream_t)stream>>>((const T*)x, ldx, rows, cols, rpb, (float*)scratch)));
  B200_LAUNCHED("colsum_partial_kernel");
  colsum_final_kernel<<<(cols + 255) / 256, 256, 0, (cudaStream_t)stream>>>((const float*)scratch, parts, cols, out);
  B200_LAUNCHED("colsum_final_kernel");
  return 0;
}

extern "C" int b200_dropout(int dtype, const void* x, void* y, long long n, float p, const unsigned long long* seed_dev,
                            unsigned long long stream_id, void* stream) {
  if (!x || !y) return fail(B200_ERR_ARG, "dropout: null pointer");
  if (n & 3) return fail(B200_ERR_UNSUPPORTED, "dropout: element count must be a multiple of 4");
  if (p > 0.f && !seed_dev) return fail(B200_ERR_ARG, "dropout: needs seed_dev");
  long long blocks = (n / 4 + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  DISPATCH_T(dtype, (dropout_kernel<T><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((const T*)x, (T*)y, n, p, seed_dev, stream_id)));
  B200_LAUNCHED("dropout_kernel");
  return 0;
}

extern "C" int b200_act_bwd(int dtype, const void* dy, const void* aux, void* out, long long n, int mode, void* stream) {
  if (!dy || !aux || !out) return fail(B200_ERR_ARG, "act_bwd: null pointer");
  if (n & 3) return fail(B200_ERR_UNSUPPORTED, "act_bwd: element count must be a multiple of 4");
  long long blocks = (n / 4 + 255) / 256;
  const long long cap = (long long)num_sms() * 8;
  if (blocks > cap) blocks = cap;
  if (blocks < 1) blocks = 1;
  DISPATCH_T(dtype, (act_bwd_kernel<T><<<(int)blocks, 256, 0, (cudaStream_t)stream>>>((const T*)dy, (const T*)aux, (T*)out, n, mode)));
  B200_LAUNCHED("act_bwd_kernel");
  return 0;
}

// ------------------------------------------------------------------------------------------------------------
// b200_colsum_static: the same column sum as b200_colsum, for callers that must not allocate device memory.
//
// Who that is: a host program that has already handed the whole device to its own pool allocator (a
// cudaMalloc from inside this library would simply fail, or fragment the pool), and a latency-sensitive loop
// that cannot afford the device-wide synchronisation cudaMalloc / cudaFree imply.  Such a caller also cannot
// hand us a scratch buffer it never planned for, so the per-block partial sums live in a fixed array that is
// part of the library image: kStaticColsumParts x kStaticColsumMaxCols floats = 1 MiB.  Every GPU that loads
// the library keeps that 1 MiB resident until the process exits, whether or not this entry point is used.
//
// `g_colsum_static_scratch` below is a __device__ symbol.  In host code that name is only a handle for the
// runtime, not a device address — passing it to a kernel directly would fault (or, where the GPU can reach host
// memory, quietly scribble over host memory).  The address the kernels receive therefore comes from
// cudaGetSymbolAddress, looked up once per device with that device current; the first call on each device does
// this set-up (it is a host-side lookup, no allocation).  Make that first call before starting any stream
// capture, so the lookup never lands inside a capture.
//
// The array is shared by every call and keeps whatever the previous call left in it.  Nothing in it has to be
// zero beforehand: each launch overwrites every partial it later reads.  Because it is shared, two calls that
// can run on the GPU at the same time (different streams, different host threads) race on it — issue them on
// one stream or order them with events.
// ------------------------------------------------------------------------------------------------------------
namespace b200 {
constexpr int kStaticColsumParts = 64;
constexpr int kStaticColsumMaxCols = 4096;
static __device__ __align__(256) float g_colsum_static_scratch[kStaticColsumParts * kStaticColsumMaxCols];

constexpr int kMaxDevices = 64;
static float* g_colsum_static_ptr[kMaxDevices] = {nullptr};
static std::mutex g_colsum_static_mu;

// device address of the static scratch on the current device (looked up once per device)
static int colsum_static_scratch(float** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return cuda_fail(e, "colsum_static: cudaGetDevice");
  if (dev < 0 || dev >= kMaxDevices) return fail(B200_ERR_UNSUPPORTED, "colsum_static: device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lock(g_colsum_static_mu);
  if (g_colsum_static_ptr[dev] == nullptr) {
    float* p = nullptr;
    e = cudaGetSymbolAddress(reinterpret_cast<void**>(&p), g_colsum_static_scratch);
    if (e != cudaSuccess) return cuda_fail(e, "colsum_static: cudaGetSymbolAddress");
    if (p == nullptr) return fail(B200_ERR_STATE, "colsum_static: symbol address is null");
    g_colsum_static_ptr[dev] = p;
  }
  *out = g_colsum_static_ptr[dev];
  return 0;
}
}  // namespace b200

extern "C" int b200_colsum_static_max_cols(void) { return b200::kStaticColsumMaxCols; }

extern "C" int b200_colsum_static(int dtype, const void* x, long long ldx, int rows, int cols, float* out, void* stream) {
  if (!x || !out) return fail(B200_ERR_ARG, "colsum_static: null pointer");
  if (rows <= 0 || cols <= 0) return fail(B200_ERR_ARG, "colsum_static: empty input (%d x %d)", rows, cols);
  if (cols & 3) return fail(B200_ERR_UNSUPPORTED, "colsum_static: cols must be a multiple of 4");
  if (cols > kStaticColsumMaxCols)
    return fail(B200_ERR_ARG, "colsum_static: cols = %d exceeds the built-in scratch (max %d columns); use b200_colsum with your own scratch",
                cols, kStaticColsumMaxCols);
  if (const int lrc = colsum_check_layout("colsum_static", dtype, x, ldx, cols)) return lrc;
  float* scratch = nullptr;
  const int rc = colsum_static_scratch(&scratch);
  if (rc) return rc;
  const int parts = rows >= 2048 ? kStaticColsumParts : (rows >= 256 ? 16 : 1);
  const int rpb = (rows + parts - 1) / parts;
  dim3 grid((cols + 127) / 128, parts);
  DISPATCH_T(dtype, (colsum_partial_kernel<T><<<grid, 256, 0, (cudaStream_t)stream>>>((const T*)x, ldx, rows, cols, rpb, scratch)));
  B200_LAUNCHED("colsum_partial_kernel(static)");
  colsum_final_kernel<<<(cols + 255) / 256, 256, 0, (cudaStream_t)stream>>>(scratch, parts, cols, out);
  B200_LAUNCHED("colsum_final_kernel(static)");
  return 0;
}
