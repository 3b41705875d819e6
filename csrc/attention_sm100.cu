// attention_sm100.cu — fused multi-head attention for head size 64 on sm_100a (tcgen05 + TMEM + TMA).
//
//   forward : O = dropout(softmax(Q K^T / sqrt(d) + mask)) V          — no [B,h,L,L] tensor in HBM
//   backward: dQ, dK, dV with the probabilities recomputed from the saved log-sum-exp
//
// Replaces hetseq/bert_modeling.py:341-376 (two batched cuBLAS GEMMs + scale + mask + Softmax + Dropout, which
// materialise B*h*L*L probabilities) and its autograd dual.
//
// One CTA = 128 threads = 128 rows of one (batch, head): thread t owns TMEM lane t, i.e. one row of the score
// tile, so softmax statistics need no cross-thread reduction.  Thread 0 issues TMA and tcgen05.mma.  Q/K/V are
// read straight out of the packed [B*L, 3H] projection through 4-D tensor maps (d, L, head, batch): rows past L
// are zero-filled by TMA, so ragged sequence lengths need no special casing.
//   S = Q K^T        : A = Q tile, B = K tile, both K-major (d contiguous), 128x128 fp32 accumulator in TMEM
//   O_j = P V        : P is written by its row owners into the (dead) K buffer as a K-major A operand, 64 keys
//                      at a time; V is the MN-major B operand (fp32: 32-byte-atom swizzle; bf16: same bytes as
//                      the K-major image)
// Dropout masks use the same (seed, stream id, element index) scheme as the stand-alone softmax kernels, so the
// fused and unfused paths are interchangeable bit-for-bit in which elements they drop.
#include "common.h"
#include "sm100_ptx.cuh"
#include "vec_utils.cuh"
#include "../include/b200seq.h"

#include <mutex>

namespace b200 {

int make_tmap_4d(CUtensorMap* out, const void* ptr, int elem_bytes, bool is_bf16, long long rows, long long cols,
                 long long ld, int box_cols, int box_rows, bool atom32, long long nb0, long long sb0, long long nb1,
                 long long sb1);

constexpr int kAttThreads = 128;
constexpr int kTile = 16384;          // one [128 rows x 128 B] swizzled tile

struct AttnArgs {
  int B, heads, L, Lp;
  float scale, p_drop;
  const long long* mask;              // [B][L] 0/1 or null
  const unsigned long long* seed_dev;
  unsigned long long stream_id;
  void* out;                          // forward: ctx [B*L][H];   backward: dqkv [B*L][3H]
  long long ld_out;
  float* lse;                         // [B][heads][L]
  float* delta;                       // [B][heads][L]  rowsum(dO * O)
  const void* o;                      // backward: forward output [B*L][H]
  const void* dout;                   // backward: dO [B*L][H]
  long long ld_o;
};

template <typename T> struct AttCfg;
template <> struct AttCfg<float> {
  static constexpr int ES = 4, UK = 8, STEPS = 8, ATOMS = 2, EPA = 32;   // k-steps per 64-wide operand, 128 B atoms, elems/atom
  static constexpr MmaKind KIND = MmaKind::TF32;
  static constexpr uint32_t FMT = 2;
  static constexpr bool SEP_MN = true;                                  // MN-major image differs from the K-major one
  static constexpr uint32_t MN_SBO = 512, MN_LT = 1;
};
template <> struct AttCfg<__nv_bfloat16> {
  static constexpr int ES = 2, UK = 16, STEPS = 4, ATOMS = 1, EPA = 64;
  static constexpr MmaKind KIND = MmaKind::F16;
  static constexpr uint32_t FMT = 1;
  static constexpr bool SEP_MN = false;
  static constexpr uint32_t MN_SBO = 1024, MN_LT = 2;
};

// K-major [rows x 64] operand: k-step j of STEPS
template <typename T>
__device__ __forceinline__ uint64_t desc_k(uint32_t base, int j) {
  using C = AttCfg<T>;
  const int per_atom = C::STEPS / C::ATOMS;
  return make_smem_desc_sw128(base + (j / per_atom) * kTile + (j % per_atom) * 32, 16, 1024);
}
// MN-major operand: rows = reduction dim (128 B each), N = 64 contiguous; k-step j covers UK rows from row0
template <typename T>
__device__ __forceinline__ uint64_t desc_mn(uint32_t base, int row0, int j) {
  using C = AttCfg<T>;
  return make_smem_desc_sw128(base + (row0 + j * C::UK) * 128, kTile, C::MN_SBO, C::MN_LT);
}

// TMA: [128 rows x 64] tile of head h / batch b starting at row r0 -> ATOMS swizzled tiles at dst
template <typename T>
__device__ __forceinline__ void load_tile(uint8_t* dst, const CUtensorMap* tm, uint64_t* bar, int r0, int h, int b) {
  using C = AttCfg<T>;
#pragma unroll
  for (int a = 0; a < C::ATOMS; ++a) tma_load_4d(dst + a * kTile, tm, bar, a * C::EPA, r0, h, b);
}

// row owner writes 64 values (one half of a probability / dS row) as a K-major A operand
template <typename T>
__device__ __forceinline__ void store_half_row(uint8_t* base, int row, const float (&v)[64]) {
  if constexpr (sizeof(T) == 4) {
#pragma unroll
    for (int a = 0; a < 2; ++a) {
      uint8_t* rp = base + a * kTile + row * 128;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const float* s = &v[a * 32 + i * 4];
        *reinterpret_cast<float4*>(rp + ((i ^ (row & 7)) << 4)) = make_float4(s[0], s[1], s[2], s[3]);
      }
    }
  } else {
    uint8_t* rp = base + row * 128;
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const float* s = &v[i * 8];
      __nv_bfloat162 p0 = __floats2bfloat162_rn(s[0], s[1]), p1 = __floats2bfloat162_rn(s[2], s[3]);
      __nv_bfloat162 p2 = __floats2bfloat162_rn(s[4], s[5]), p3 = __floats2bfloat162_rn(s[6], s[7]);
      uint4 u;
      u.x = *reinterpret_cast<uint32_t*>(&p0); u.y = *reinterpret_cast<uint32_t*>(&p1);
      u.z = *reinterpret_cast<uint32_t*>(&p2); u.w = *reinterpret_cast<uint32_t*>(&p3);
      *reinterpret_cast<uint4*>(rp + ((i ^ (row & 7)) << 4)) = u;
    }
  }
}

__device__ __forceinline__ void ld32(uint32_t taddr, float (&v)[32]) {
  uint32_t r[32];
  tmem_ld_32x32(taddr, r);
  tmem_ld_wait();
#pragma unroll
  for (int j = 0; j < 32; ++j) v[j] = __uint_as_float(r[j]);
}

template <typename T>
__device__ __forceinline__ void load_row64(const T* p, float (&v)[64]) {
#pragma unroll
  for (int i = 0; i < 16; ++i) {
    const f4 t = load4<T>(p + i * 4);
    v[i * 4] = t.x; v[i * 4 + 1] = t.y; v[i * 4 + 2] = t.z; v[i * 4 + 3] = t.w;
  }
}
template <typename T>
__device__ __forceinline__ void store_row64(T* p, const float (&v)[64]) {
#pragma unroll
  for (int i = 0; i < 16; ++i) store4<T>(p + i * 4, f4{v[i * 4], v[i * 4 + 1], v[i * 4 + 2], v[i * 4 + 3]});
}

// =====================================================================================================
// backward, part 1: dQ (and delta = rowsum(dO * O)).  Thread = query row.  grid = (q blocks, heads, B)
// =====================================================================================================
// smem tiles: Q, dO (A operands, loaded once), K (B of S), V (B of dP; later dS halves), K in MN-major form (B of dQ;
// for bf16 the K-major image already is that form).  TMEM: S [0,128), dP [128,256), dQ [256,320).
template <typename T>
__global__ void __launch_bounds__(kAttThreads)
attn_bwd_dq_kernel(const __grid_constant__ CUtensorMap tmQK, const __grid_constant__ CUtensorMap tmMN,
                   const __grid_constant__ CUtensorMap tmdO, const AttnArgs a) {
  using C = AttCfg<T>;
  constexpr int OPB = C::ATOMS * kTile;
  constexpr int NT = C::SEP_MN ? 5 : 4;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sdO = smem + OPB;
  uint8_t* sK = smem + 2 * OPB;
  uint8_t* sV = smem + 3 * OPB;                  // dS halves once dP is complete
  uint8_t* sKmn = C::SEP_MN ? smem + 4 * OPB : sK;
  float* sMask = reinterpret_cast<float*>(smem + NT * OPB);
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + NT * OPB + 512);
  uint64_t *bar_q = bars, *bar_kv = bars + 1, *bar_s = bars + 2, *bar_pv = bars + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int t = threadIdx.x, warp = t >> 5;
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = blockIdx.x * 128;
  const int qrow = q0 + t;
  const bool qlive = qrow < a.L;

  if (t == 0) {
    tma_prefetch_desc(&tmQK); tma_prefetch_desc(&tmMN); tma_prefetch_desc(&tmdO);
    mbar_init(bar_q, 1); mbar_init(bar_kv, 1); mbar_init(bar_s, 1); mbar_init(bar_pv, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem + ((uint32_t)(warp * 32) << 16);

  constexpr uint32_t IDESC_S = make_idesc(C::FMT, false, false, 128, 128);
  constexpr uint32_t IDESC_Q = make_idesc(C::FMT, false, true, 128, 64);

  if (t == 0) {
    mbar_arrive_expect_tx(bar_q, 2 * OPB);
    load_tile<T>(sQ, &tmQK, bar_q, q0, h, b);
    load_tile<T>(sdO, &tmdO, bar_q, q0, h, b);
  }

  // delta = rowsum(dO * O) for this query row; also published for the dK/dV kernel
  float delta = 0.f, lse = 0.f;
  if (qlive) {
    const T* orow = reinterpret_cast<const T*>(a.o) + ((long long)b * a.L + qrow) * a.ld_o + h * 64;
    const T* drow = reinterpret_cast<const T*>(a.dout) + ((long long)b * a.L + qrow) * a.ld_o + h * 64;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
      const f4 x = load4<T>(orow + i * 4), y = load4<T>(drow + i * 4);
      delta += x.x * y.x + x.y * y.y + x.z * y.z + x.w * y.w;
    }
    const long long si = ((long long)b * a.heads + h) * a.L + qrow;
    a.delta[si] = delta;
    lse = a.lse[si];
  }

  const Philox ph(a.seed_dev ? *a.seed_dev : 0ull);
  const uint32_t thr = dropout_threshold(a.p_drop);
  const float ik = a.p_drop > 0.f ? 1.f / (1.f - a.p_drop) : 1.f;
  const long long rowg = (((long long)b * a.heads + h) * a.L + qrow) * a.Lp;

  const int nkv = (a.L + 127) / 128;
  for (int j = 0; j < nkv; ++j) {
    const int k0 = j * 128;
    if (t == 0) {
      mbar_arrive_expect_tx(bar_kv, (C::SEP_MN ? 3 : 2) * OPB);
      load_tile<T>(sK, &tmQK, bar_kv, k0, a.heads + h, b);
      load_tile<T>(sV, &tmQK, bar_kv, k0, 2 * a.heads + h, b);
      if constexpr (C::SEP_MN) load_tile<T>(sKmn, &tmMN, bar_kv, k0, a.heads + h, b);
    }
    {
      const int k = k0 + t;
      float mv = -3.0e38f;
      if (k < a.L) mv = a.mask ? (1.f - (float)a.mask[(long long)b * a.L + k]) * -10000.f : 0.f;
      sMask[t] = mv;
    }
    if (t == 0) {
      if (j == 0) mbar_wait(bar_q, 0);
      mbar_wait(bar_kv, j & 1);
      tcgen05_fence_after();
      const uint32_t qa = smem_u32(sQ), da = smem_u32(sdO), ka = smem_u32(sK), va = smem_u32(sV);
#pragma unroll
      for (int s = 0; s < C::STEPS; ++s)
        tcgen05_mma_ss<C::KIND>(tmem, desc_k<T>(qa, s), desc_k<T>(ka, s), IDESC_S, s > 0 ? 1u : 0u);
#pragma unroll
      for (int s = 0; s < C::STEPS; ++s)
        tcgen05_mma_ss<C::KIND>(tmem + 128, desc_k<T>(da, s), desc_k<T>(va, s), IDESC_S, s > 0 ? 1u : 0u);
      tcgen05_commit(bar_s);
    }
    __syncthreads();
    mbar_wait(bar_s, j & 1);
    tcgen05_fence_after();

#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      float ds[64];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float sv[32], dp[32];
        ld32(tS + half * 64 + c * 32, sv);
        ld32(tS + 128 + half * 64 + c * 32, dp);
        if (a.p_drop > 0.f) {
#pragma unroll
          for (int i = 0; i < 32; i += 4) {
            const f4 d = dropout_scale4(ph, (unsigned long long)(rowg + k0 + half * 64 + c * 32 + i), a.stream_id, thr, ik);
            dp[i] *= d.x; dp[i + 1] *= d.y; dp[i + 2] *= d.z; dp[i + 3] *= d.w;
          }
        }
#pragma unroll
        for (int i = 0; i < 32; ++i) {
          const float mv = sMask[half * 64 + c * 32 + i];
          const float p = (mv > -1.0e38f && qlive) ? __expf(sv[i] * a.scale + mv - lse) : 0.f;
          ds[c * 32 + i] = p * (dp[i] - delta);
        }
      }
      store_half_row<T>(sV, t, ds);
      fence_proxy_async_smem();
      tcgen05_fence_before();
      __syncthreads();
      if (t == 0) {
        tcgen05_fence_after();
        const uint32_t sa = smem_u32(sV), kb = smem_u32(sKmn);
#pragma unroll
        for (int s = 0; s < C::STEPS; ++s)
          tcgen05_mma_ss<C::KIND>(tmem + 256, desc_k<T>(sa, s), desc_mn<T>(kb, half * 64, s), IDESC_Q,
                                  (j > 0 || half > 0 || s > 0) ? 1u : 0u);
        tcgen05_commit(bar_pv);
      }
      mbar_wait(bar_pv, (2 * j + half) & 1);
      tcgen05_fence_after();
    }
    tcgen05_fence_before();
    __syncthreads();
  }

  {
    float dq[64];      // TMEM loads are warp-collective: issue them on every lane, store only live rows
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      float v[32];
      ld32(tS + 256 + c * 32, v);
#pragma unroll
      for (int i = 0; i < 32; ++i) dq[c * 32 + i] = v[i] * a.scale;
    }
    if (qlive) {
      T* dst = reinterpret_cast<T*>(a.out) + ((long long)b * a.L + qrow) * a.ld_out + h * 64;    // Q section of dqkv
      store_row64<T>(dst, dq);
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

// =====================================================================================================
// backward, part 2: dK, dV.  Thread = key row.  grid = (kv blocks, heads, B)
// =====================================================================================================
// S^T = K Q^T and dP^T = V dO^T put keys on TMEM lanes, so the key's owner forms Pd^T and dS^T rows directly.
// smem tiles: K, V (A operands, loaded once), Q, dO in K-major form (B of S^T / dP^T), Q, dO in MN-major form
// (B of dK = dS^T Q, dV = Pd^T dO), and the two 64-query halves of Pd^T / dS^T (A operands).
//   fp32: the halves overlay the dead K-major Q / dO tiles; bf16: K-major == MN-major image, halves get own tiles.
// TMEM: S^T [0,128), dP^T [128,256), dV [256,320), dK [320,384).
template <typename T>
__global__ void __launch_bounds__(kAttThreads)
attn_bwd_dkv_kernel(const __grid_constant__ CUtensorMap tmQK, const __grid_constant__ CUtensorMap tmMN,
                    const __grid_constant__ CUtensorMap tmdO, const __grid_constant__ CUtensorMap tmdOmn,
                    const AttnArgs a) {
  using C = AttCfg<T>;
  constexpr int OPB = C::ATOMS * kTile;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sK = smem;
  uint8_t* sV = smem + OPB;
  uint8_t* sQ = smem + 2 * OPB;
  uint8_t* sdO = smem + 3 * OPB;
  uint8_t* sQmn = C::SEP_MN ? smem + 4 * OPB : sQ;
  uint8_t* sdOmn = C::SEP_MN ? smem + 5 * OPB : sdO;
  uint8_t* sX = C::SEP_MN ? sQ : smem + 4 * OPB;      // Pd^T half
  uint8_t* sY = C::SEP_MN ? sdO : smem + 5 * OPB;     // dS^T half
  float* sLse = reinterpret_cast<float*>(smem + 6 * OPB);
  float* sDelta = sLse + 128;
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem + 6 * OPB + 1024);
  uint64_t *bar_kv = bars, *bar_q = bars + 1, *bar_s = bars + 2, *bar_pv = bars + 3;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 4);

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int h = blockIdx.y, b = blockIdx.z;
  const int k0 = blockIdx.x * 128;
  const int krow = k0 + t;
  const bool klive = krow < a.L;

  if (t == 0) {
    tma_prefetch_desc(&tmQK); tma_prefetch_desc(&tmMN); tma_prefetch_desc(&tmdO); tma_prefetch_desc(&tmdOmn);
    mbar_init(bar_kv, 1); mbar_init(bar_q, 1); mbar_init(bar_s, 1); mbar_init(bar_pv, 1);
    fence_barrier_init();
  }
  if (warp == 0) tmem_alloc<512>(tmem_slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem + ((uint32_t)(warp * 32) << 16);

  constexpr uint32_t IDESC_S = make_idesc(C::FMT, false, false, 128, 128);
  constexpr uint32_t IDESC_G = make_idesc(C::FMT, false, true, 128, 64);

  if (t == 0) {
    mbar_arrive_expect_tx(bar_kv, 2 * OPB);
    load_tile<T>(sK, &tmQK, bar_kv, k0, a.heads + h, b);
    load_tile<T>(sV, &tmQK, bar_kv, k0, 2 * a.heads + h, b);
  }
  float mv = -3.0e38f;
  if (klive) mv = a.mask ? (1.f - (float)a.mask[(long long)b * a.L + krow]) * -10000.f : 0.f;

  const Philox ph(a.seed_dev ? *a.seed_dev : 0ull);
  const uint32_t thr = dropout_threshold(a.p_drop);
  const float ik = a.p_drop > 0.f ? 1.f / (1.f - a.p_drop) : 1.f;
  const long long bh = (long long)b * a.heads + h;
  const int r4 = lane & 3;

  const int nq = (a.L + 127) / 128;
  for (int i = 0; i < nq; ++i) {
    const int q0 = i * 128;
    if (t == 0) {
      mbar_arrive_expect_tx(bar_q, (C::SEP_MN ? 4 : 2) * OPB);
      load_tile<T>(sQ, &tmQK, bar_q, q0, h, b);
      load_tile<T>(sdO, &tmdO, bar_q, q0, h, b);
      if constexpr (C::SEP_MN) {
        load_tile<T>(sQmn, &tmMN, bar_q, q0, h, b);
        load_tile<T>(sdOmn, &tmdOmn, bar_q, q0, h, b);
      }
    }
    {
      const int q = q0 + t;
      sLse[t] = q < a.L ? a.lse[bh * a.L + q] : 3.0e38f;      // +huge -> exp(...) = 0 for rows past L
      sDelta[t] = q < a.L ? a.delta[bh * a.L + q] : 0.f;
    }
    if (t == 0) {
      if (i == 0) mbar_wait(bar_kv, 0);
      mbar_wait(bar_q, i & 1);
      tcgen05_fence_after();
      const uint32_t ka = smem_u32(sK), va = smem_u32(sV), qa = smem_u32(sQ), da = smem_u32(sdO);
#pragma unroll
      for (int s = 0; s < C::STEPS; ++s)
        tcgen05_mma_ss<C::KIND>(tmem, desc_k<T>(ka, s), desc_k<T>(qa, s), IDESC_S, s > 0 ? 1u : 0u);
#pragma unroll
      for (int s = 0; s < C::STEPS; ++s)
        tcgen05_mma_ss<C::KIND>(tmem + 128, desc_k<T>(va, s), desc_k<T>(da, s), IDESC_S, s > 0 ? 1u : 0u);
      tcgen05_commit(bar_s);
    }
    __syncthreads();
    mbar_wait(bar_s, i & 1);
    tcgen05_fence_after();

#pragma unroll 1
    for (int half = 0; half < 2; ++half) {
      float pd[64], ds[64];
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        float sv[32], dp[32];
        ld32(tS + half * 64 + c * 32, sv);
        ld32(tS + 128 + half * 64 + c * 32, dp);
#pragma unroll
        for (int g = 0; g < 32; g += 4) {
          // dropout factors of (query qq..qq+3, this key): each lane of a 4-lane group draws the 4 keys of ONE of the
          // queries (same counter the forward used), then the group transposes them with shuffles
          float dm[4] = {1.f, 1.f, 1.f, 1.f};
          if (a.p_drop > 0.f) {
            const int qq = q0 + half * 64 + c * 32 + g;
            const long long idx = ((bh * a.L + qq + r4) * a.Lp + (krow & ~3));
            const f4 mine = dropout_scale4(ph, (unsigned long long)idx, a.stream_id, thr, ik);
#pragma unroll
            for (int s = 0; s < 4; ++s) {
              const int sel = (r4 - s) & 3;
              const float send = sel == 0 ? mine.x : (sel == 1 ? mine.y : (sel == 2 ? mine.z : mine.w));
              const int src = (lane & ~3) | ((r4 + s) & 3);
              const float got = __shfl_sync(0xffffffffu, send, src);
              const int dst = (r4 + s) & 3;
              if (dst == 0) dm[0] = got; else if (dst == 1) dm[1] = got; else if (dst == 2) dm[2] = got; else dm[3] = got;
            }
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int qi = half * 64 + c * 32 + g + e;
            const float p = klive ? __expf(sv[g + e] * a.scale + mv - sLse[qi]) : 0.f;
            pd[c * 32 + g + e] = p * dm[e];
            ds[c * 32 + g + e] = p * (dp[g + e] * dm[e] - sDelta[qi]);
          }
        }
      }
      store_half_row<T>(sX, t, pd);
      store_half_row<T>(sY, t, ds);
      fence_proxy_async_smem();
      tcgen05_fence_before();
      __syncthreads();
      if (t == 0) {
        tcgen05_fence_after();
        const uint32_t xa = smem_u32(sX), ya = smem_u32(sY), qm = smem_u32(sQmn), dm_ = smem_u32(sdOmn);
#pragma unroll
        for (int s = 0; s < C::STEPS; ++s)
          tcgen05_mma_ss<C::KIND>(tmem + 256, desc_k<T>(xa, s), desc_mn<T>(dm_, half * 64, s), IDESC_G,
                                  (i > 0 || half > 0 || s > 0) ? 1u : 0u);
#pragma unroll
        for (int s = 0; s < C::STEPS; ++s)
          tcgen05_mma_ss<C::KIND>(tmem + 320, desc_k<T>(ya, s), desc_mn<T>(qm, half * 64, s), IDESC_G,
                                  (i > 0 || half > 0 || s > 0) ? 1u : 0u);
        tcgen05_commit(bar_pv);
      }
      mbar_wait(bar_pv, (2 * i + half) & 1);
      tcgen05_fence_after();
    }
    tcgen05_fence_before();
    __syncthreads();
  }

  {
    float dv[64], dk[64];
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      float v[32];
      ld32(tS + 256 + c * 32, v);
#pragma unroll
      for (int e = 0; e < 32; ++e) dv[c * 32 + e] = v[e];
      ld32(tS + 320 + c * 32, v);
#pragma unroll
      for (int e = 0; e < 32; ++e) dk[c * 32 + e] = v[e] * a.scale;
    }
    if (klive) {
      T* base = reinterpret_cast<T*>(a.out) + ((long long)b * a.L + krow) * a.ld_out + h * 64;
      store_row64<T>(base + (long long)a.heads * 64, dk);          // K section of dqkv
      store_row64<T>(base + (long long)2 * a.heads * 64, dv);      // V section
    }
  }
  tcgen05_fence_before();
  __syncthreads();
  if (warp == 0) {
    tcgen05_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

}  // namespace b200

#include "attention_wide.cuh"

namespace b200 {

// =====================================================================================================
// host side
// =====================================================================================================
template <typename T> constexpr int fwd_smem() { return 5 * AttCfg<T>::ATOMS * kTile + 128 * 4 + 512 * 4 + 128 + 1024; }
// the fp32 single-kernel backward fills the SM: 224 KB of tiles + 1.6 KB of row data; no room for alignment slack, the
// kernel traps if the dynamic window is not 1 KB aligned (it is: no static shared memory in that kernel)
template <typename T> constexpr int blk_smem() {
  return sizeof(T) == 4 ? 7 * AttCfg<T>::ATOMS * kTile + 2 * 512 + 128 : 8 * AttCfg<T>::ATOMS * kTile + 2 * 512 + 128 + 1024;
}
template <typename T> constexpr int dq_smem() { return (AttCfg<T>::SEP_MN ? 5 : 4) * AttCfg<T>::ATOMS * kTile + 512 + 128 + 1024; }
template <typename T> constexpr int dkv_smem() { return 6 * AttCfg<T>::ATOMS * kTile + 1024 + 128 + 1024; }

// tensor map over rows of a packed [B*L, nsec*heads*64] matrix, seen as (d=64, L, nsec*heads, B)
template <typename T>
static int head_map(CUtensorMap* tm, const void* p, int B, int heads_total, int L, long long ld, bool mn) {
  constexpr bool f32 = sizeof(T) == 4;
  return make_tmap_4d(tm, p, sizeof(T), !f32, L, 64, ld, f32 ? 32 : 64, 128, mn && f32, heads_total, 64, B, (long long)L * ld);
}

template <typename T>
static int attn_fwd_impl(const b200_attn_args* g, cudaStream_t st) {
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFuncSetAttribute(attn_fwd_wide_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, fwd_smem<T>());
  });
  CUtensorMap tq, tv, to;
  if (int e = head_map<T>(&tq, g->qkv, g->B, 3 * g->heads, g->L, g->ld_qkv, false)) return e;
  if (int e = head_map<T>(&tv, g->qkv, g->B, 3 * g->heads, g->L, g->ld_qkv, true)) return e;
  if (int e = head_map<T>(&to, g->out, g->B, g->heads, g->L, g->ld_out, false)) return e;
  AttnArgs a{};
  a.B = g->B; a.heads = g->heads; a.L = g->L; a.Lp = g->Lp;
  a.scale = g->scale; a.p_drop = g->p_drop; a.mask = g->attn_mask; a.seed_dev = g->seed_dev; a.stream_id = g->stream_id;
  a.out = g->out; a.ld_out = g->ld_out; a.lse = g->lse;
  dim3 grid((g->L + 127) / 128, g->heads, g->B);
  B200_CUDA(launch_pdl(attn_fwd_wide_kernel<T>, grid, dim3(kWideThreads), fwd_smem<T>(), st, 1, tq, tv, to, a));
  B200_LAUNCHED("attn_fwd_wide_kernel");
  return 0;
}

template <typename T>
static int attn_bwd_impl(const b200_attn_args* g, cudaStream_t st) {
  static std::once_flag once;
  std::call_once(once, [] {
    cudaFuncSetAttribute(attn_bwd_dq_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, dq_smem<T>());
    cudaFuncSetAttribute(attn_bwd_dkv_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, dkv_smem<T>());
    cudaFuncSetAttribute(attn_bwd_block_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, blk_smem<T>());
  });
  CUtensorMap tq, tmn, tdo, tdomn, tout;
  if (int e = head_map<T>(&tq, g->qkv, g->B, 3 * g->heads, g->L, g->ld_qkv, false)) return e;
  if (int e = head_map<T>(&tmn, g->qkv, g->B, 3 * g->heads, g->L, g->ld_qkv, true)) return e;
  if (int e = head_map<T>(&tdo, g->dout, g->B, g->heads, g->L, g->ld_o, false)) return e;
  if (int e = head_map<T>(&tdomn, g->dout, g->B, g->heads, g->L, g->ld_o, true)) return e;
  AttnArgs a{};
  a.B = g->B; a.heads = g->heads; a.L = g->L; a.Lp = g->Lp;
  a.scale = g->scale; a.p_drop = g->p_drop; a.mask = g->attn_mask; a.seed_dev = g->seed_dev; a.stream_id = g->stream_id;
  a.out = g->out; a.ld_out = g->ld_out; a.lse = g->lse; a.delta = g->delta;
  a.o = g->o; a.dout = g->dout; a.ld_o = g->ld_o;
  if (g->L <= 128) {       // one kernel computes all three gradients of a (batch, head)
    if (int e = head_map<T>(&tout, g->out, g->B, 3 * g->heads, g->L, g->ld_out, false)) return e;
    B200_CUDA(launch_pdl(attn_bwd_block_kernel<T>, dim3(g->heads, g->B), dim3(kWideThreads), blk_smem<T>(), st, 1, tq, tmn, tdo,
                         tdomn, tout, a));
    B200_LAUNCHED("attn_bwd_block_kernel");
    return 0;
  }
  dim3 grid((g->L + 127) / 128, g->heads, g->B);
  attn_bwd_dq_kernel<T><<<grid, kAttThreads, dq_smem<T>(), st>>>(tq, tmn, tdo, a);
  B200_LAUNCHED("attn_bwd_dq_kernel");
  attn_bwd_dkv_kernel<T><<<grid, kAttThreads, dkv_smem<T>(), st>>>(tq, tmn, tdo, tdomn, a);
  B200_LAUNCHED("attn_bwd_dkv_kernel");
  return 0;
}

static int attn_check(const b200_attn_args* g, bool bwd) {
  if (!g || !g->qkv || !g->out || !g->lse) return fail(B200_ERR_ARG, "attn: null argument");
  if (g->B < 1 || g->heads < 1 || g->L < 1) return fail(B200_ERR_ARG, "attn: bad shape");
  if (g->Lp < g->L || (g->Lp & 3)) return fail(B200_ERR_ARG, "attn: Lp must be >= L and a multiple of 4");
  if (g->dtype != 0 && g->dtype != 1) return fail(B200_ERR_UNSUPPORTED, "attn: dtype 0 (fp32 via tf32) or 1 (bf16) only");
  if (g->p_drop > 0.f && !g->seed_dev) return fail(B200_ERR_ARG, "attn: dropout needs seed_dev");
  if (bwd && (!g->dout || !g->delta || (g->L > 128 && !g->o))) return fail(B200_ERR_ARG, "attn backward: null argument");
  return 0;
}

}  // namespace b200

extern "C" int b200_attn_fwd(const b200_attn_args* g, void* stream) {
  using namespace b200;
  if (int e = attn_check(g, false)) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return g->dtype == 0 ? attn_fwd_impl<float>(g, st) : attn_fwd_impl<__nv_bfloat16>(g, st);
}

extern "C" int b200_attn_bwd(const b200_attn_args* g, void* stream) {
  using namespace b200;
  if (int e = attn_check(g, true)) return e;
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  return g->dtype == 0 ? attn_bwd_impl<float>(g, st) : attn_bwd_impl<__nv_bfloat16>(g, st);
}
