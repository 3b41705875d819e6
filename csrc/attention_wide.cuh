// attention_wide.cuh — the 512-thread attention kernels (included by attention_sm100.cu).
//
// Sixteen warps share one 128-row tile: warp w owns TMEM lanes 32*(w&3).. (the hardware's lane quadrant rule) and the
// 32-column slice (w>>2) of every 128-column score tile, so the exp / Philox / convert work of a tile is spread over
// 512 threads instead of 128.  Row statistics cross the four column slices through a [4][128] shared array.
#pragma once

namespace b200 {

constexpr int kWideThreads = 512;

// thread (row, slice cq) writes its 32 probabilities into the K-major [128 x 128] A-operand image
template <typename T>
__device__ __forceinline__ void store_slice_row(uint8_t* base, int row, int cq, const float (&v)[32]) {
  if constexpr (sizeof(T) == 4) {
    uint8_t* rp = base + cq * kTile + row * 128;                 // atom cq: 32 fp32 = one 128 B row
#pragma unroll
    for (int i = 0; i < 8; ++i)
      *reinterpret_cast<float4*>(rp + ((i ^ (row & 7)) << 4)) = make_float4(v[i * 4], v[i * 4 + 1], v[i * 4 + 2], v[i * 4 + 3]);
  } else {
    uint8_t* rp = base + (cq >> 1) * kTile + row * 128;          // atom cq/2: 64 bf16; this slice is half a row
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const float* s = &v[i * 8];
      __nv_bfloat162 p0 = __floats2bfloat162_rn(s[0], s[1]), p1 = __floats2bfloat162_rn(s[2], s[3]);
      __nv_bfloat162 p2 = __floats2bfloat162_rn(s[4], s[5]), p3 = __floats2bfloat162_rn(s[6], s[7]);
      uint4 u;
      u.x = *reinterpret_cast<uint32_t*>(&p0); u.y = *reinterpret_cast<uint32_t*>(&p1);
      u.z = *reinterpret_cast<uint32_t*>(&p2); u.w = *reinterpret_cast<uint32_t*>(&p3);
      const int chunk = (cq & 1) * 4 + i;
      *reinterpret_cast<uint4*>(rp + ((chunk ^ (row & 7)) << 4)) = u;
    }
  }
}

// fp32 only: the same 32 values written transposed, image [128 key rows x 128 queries] K-major (atom = row / 32).
// A warp's 32 lanes are 32 consecutive queries -> one 128 B row per store instruction, conflict-free.
__device__ __forceinline__ void store_slice_transposed(uint8_t* base, int row, int cq, const float (&v)[32]) {
  uint8_t* ap = base + (row >> 5) * kTile + ((row & 31) & 3) * 4;
  const int chunk = (row & 31) >> 2;
#pragma unroll
  for (int e = 0; e < 32; ++e) {
    const int k = cq * 32 + e;
    *reinterpret_cast<float*>(ap + k * 128 + ((chunk ^ (k & 7)) << 4)) = v[e];
  }
}

__device__ __forceinline__ void ld16(uint32_t taddr, float (&v)[16]) {
  uint32_t r[16];
  tmem_ld_32x16(taddr, r);
  tmem_ld_wait();
#pragma unroll
  for (int j = 0; j < 16; ++j) v[j] = __uint_as_float(r[j]);
}

template <typename T>
__device__ __forceinline__ void store_row16(T* p, const float (&v)[16]) {
#pragma unroll
  for (int i = 0; i < 4; ++i) store4<T>(p + i * 4, f4{v[i * 4], v[i * 4 + 1], v[i * 4 + 2], v[i * 4 + 3]});
}

// ---- generator forms of the slice stores: `get(i)` yields value i of the thread's 32, so callers need not
// materialise a second 32-register array --------------------------------------------------------------------------
template <typename T, typename F>
__device__ __forceinline__ void store_slice_row_gen(uint8_t* base, int row, int cq, F get) {
  if constexpr (sizeof(T) == 4) {
    uint8_t* rp = base + cq * kTile + row * 128;
#pragma unroll
    for (int i = 0; i < 8; ++i)
      *reinterpret_cast<float4*>(rp + ((i ^ (row & 7)) << 4)) = make_float4(get(i * 4), get(i * 4 + 1), get(i * 4 + 2), get(i * 4 + 3));
  } else {
    uint8_t* rp = base + (cq >> 1) * kTile + row * 128;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      __nv_bfloat162 p0 = __floats2bfloat162_rn(get(i * 8), get(i * 8 + 1)), p1 = __floats2bfloat162_rn(get(i * 8 + 2), get(i * 8 + 3));
      __nv_bfloat162 p2 = __floats2bfloat162_rn(get(i * 8 + 4), get(i * 8 + 5)), p3 = __floats2bfloat162_rn(get(i * 8 + 6), get(i * 8 + 7));
      uint4 u;
      u.x = *reinterpret_cast<uint32_t*>(&p0); u.y = *reinterpret_cast<uint32_t*>(&p1);
      u.z = *reinterpret_cast<uint32_t*>(&p2); u.w = *reinterpret_cast<uint32_t*>(&p3);
      const int chunk = (cq & 1) * 4 + i;
      *reinterpret_cast<uint4*>(rp + ((chunk ^ (row & 7)) << 4)) = u;
    }
  }
}
template <typename F>
__device__ __forceinline__ void store_slice_transposed_gen(uint8_t* base, int row, int cq, F get) {
  uint8_t* ap = base + (row >> 5) * kTile + ((row & 31) & 3) * 4;
  const int chunk = (row & 31) >> 2;
#pragma unroll
  for (int e = 0; e < 32; ++e) {
    const int k = cq * 32 + e;
    *reinterpret_cast<float*>(ap + k * 128 + ((chunk ^ (k & 7)) << 4)) = get(e);
  }
}

// thread (row, cq) puts its 16 output columns into a [128 x 64] SW128 staging tile that a TMA store then drains
template <typename T>
__device__ __forceinline__ void stage_row16(uint8_t* tile, int row, int cq, const float (&v)[16]) {
  if constexpr (sizeof(T) == 4) {
    uint8_t* rp = tile + (cq >> 1) * kTile + row * 128;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int chunk = (cq & 1) * 4 + i;
      *reinterpret_cast<float4*>(rp + ((chunk ^ (row & 7)) << 4)) = make_float4(v[i * 4], v[i * 4 + 1], v[i * 4 + 2], v[i * 4 + 3]);
    }
  } else {
    uint8_t* rp = tile + row * 128;
#pragma unroll
    for (int i = 0; i < 2; ++i) {
      const float* s = &v[i * 8];
      __nv_bfloat162 p0 = __floats2bfloat162_rn(s[0], s[1]), p1 = __floats2bfloat162_rn(s[2], s[3]);
      __nv_bfloat162 p2 = __floats2bfloat162_rn(s[4], s[5]), p3 = __floats2bfloat162_rn(s[6], s[7]);
      uint4 u;
      u.x = *reinterpret_cast<uint32_t*>(&p0); u.y = *reinterpret_cast<uint32_t*>(&p1);
      u.z = *reinterpret_cast<uint32_t*>(&p2); u.w = *reinterpret_cast<uint32_t*>(&p3);
      const int chunk = cq * 2 + i;
      *reinterpret_cast<uint4*>(rp + ((chunk ^ (row & 7)) << 4)) = u;
    }
  }
}
template <typename T>
__device__ __forceinline__ void store_tile(const CUtensorMap* tm, const uint8_t* tile, int r0, int h, int b) {
  using C = AttCfg<T>;
#pragma unroll
  for (int a = 0; a < C::ATOMS; ++a) tma_store_4d(tm, tile + a * kTile, a * C::EPA, r0, h, b);
}

// =====================================================================================================
// forward, 512 threads.  grid = (q blocks, heads, B)
// =====================================================================================================
// smem: Q | K | V(mn) | P[128 x 128] | mask[128] | red[4][128] | barriers.   TMEM: S [0,128), O_blk [128,192)
template <typename T>
__global__ void __launch_bounds__(kWideThreads, 1)
attn_fwd_wide_kernel(const __grid_constant__ CUtensorMap tmQK, const __grid_constant__ CUtensorMap tmVmn,
                     const __grid_constant__ CUtensorMap tmOut, const AttnArgs a) {
  using C = AttCfg<T>;
  constexpr int OPB = C::ATOMS * kTile;
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* sQ = smem;
  uint8_t* sK = smem + OPB;
  uint8_t* sV = smem + 2 * OPB;
  uint8_t* sP = smem + 3 * OPB;                                   // 2*OPB bytes
  float* sMask = reinterpret_cast<float*>(smem + 5 * OPB);
  float* sRed = sMask + 128;                                      // [4][128]
  uint64_t* bars = reinterpret_cast<uint64_t*>(sRed + 512);
  uint64_t *bar_q = bars, *bar_k = bars + 1, *bar_v = bars + 2, *bar_s = bars + 3, *bar_pv = bars + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 5);

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int lq = warp & 3, cq = warp >> 2;
  const int row = lq * 32 + lane;
  const int h = blockIdx.y, b = blockIdx.z;
  const int q0 = blockIdx.x * 128;
  const int qrow = q0 + row;
  const bool qlive = qrow < a.L;
  pdl_trigger();

  if (t == 0) {
    tma_prefetch_desc(&tmQK);
    tma_prefetch_desc(&tmVmn);
    tma_prefetch_desc(&tmOut);
    mbar_init(bar_q, 1); mbar_init(bar_k, 1); mbar_init(bar_v, 1); mbar_init(bar_s, 1); mbar_init(bar_pv, 1);
    fence_barrier_init();
    pdl_wait();                                   // qkv comes from the previous kernel
    mbar_arrive_expect_tx(bar_q, OPB);
    load_tile<T>(sQ, &tmQK, bar_q, q0, h, b);
  }
  if (warp == 1) tmem_alloc<256>(tmem_slot);
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  pdl_wait();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tS = tmem + ((uint32_t)(lq * 32) << 16) + cq * 32;
  const uint32_t tO = tmem + ((uint32_t)(lq * 32) << 16) + 128 + cq * 16;

  constexpr uint32_t IDESC_S = make_idesc(C::FMT, false, false, 128, 128);
  constexpr uint32_t IDESC_O = make_idesc(C::FMT, false, true, 128, 64);

  const Philox ph(a.seed_dev ? *a.seed_dev : 0ull);
  const uint32_t thr = dropout_threshold(a.p_drop);
  const float ik = a.p_drop > 0.f ? 1.f / (1.f - a.p_drop) : 1.f;
  const long long rowg = (((long long)b * a.heads + h) * a.L + qrow) * a.Lp;

  float o_acc[16];
#pragma unroll
  for (int i = 0; i < 16; ++i) o_acc[i] = 0.f;
  float m_run = -3.0e38f, l_part = 0.f;

  const int nkv = (a.L + 127) / 128;
  for (int j = 0; j < nkv; ++j) {
    const int k0 = j * 128;
    if (t == 0) {
      mbar_arrive_expect_tx(bar_k, OPB);
      load_tile<T>(sK, &tmQK, bar_k, k0, a.heads + h, b);
      mbar_arrive_expect_tx(bar_v, OPB);
      load_tile<T>(sV, &tmVmn, bar_v, k0, 2 * a.heads + h, b);
    }
    if (t >= 128 && t < 256) {
      const int k = k0 + t - 128;
      float mv = -3.0e38f;
      if (k < a.L) mv = a.mask ? (1.f - (float)a.mask[(long long)b * a.L + k]) * -10000.f : 0.f;
      sMask[t - 128] = mv;
    }
    if (t == 0) {
      if (j == 0) mbar_wait(bar_q, 0);
      mbar_wait(bar_k, j & 1);
      tcgen05_fence_after();
      const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK);
#pragma unroll
      for (int s = 0; s < C::STEPS; ++s)
        tcgen05_mma_ss<C::KIND>(tmem, desc_k<T>(qa, s), desc_k<T>(ka, s), IDESC_S, s > 0 ? 1u : 0u);
      tcgen05_commit(bar_s);
    }
    __syncthreads();
    mbar_wait(bar_s, j & 1);
    tcgen05_fence_after();

    float p[32];
    ld32(tS, p);
    float m_loc = -3.0e38f;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float mv = sMask[cq * 32 + i];
      p[i] = mv > -1.0e38f ? p[i] * a.scale + mv : -3.0e38f;
      m_loc = fmaxf(m_loc, p[i]);
    }
    sRed[cq * 128 + row] = m_loc;
    __syncthreads();
    const float m_new = fmaxf(fmaxf(m_run, fmaxf(sRed[row], sRed[128 + row])), fmaxf(sRed[256 + row], sRed[384 + row]));
    const float alpha = __expf(m_run - m_new);
    float l_loc = 0.f;
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      p[i] = p[i] > -1.0e38f ? __expf(p[i] - m_new) : 0.f;
      l_loc += p[i];
    }
    l_part = l_part * alpha + l_loc;
    m_run = m_new;
    if (a.p_drop > 0.f) {
#pragma unroll
      for (int i = 0; i < 32; i += 4) {
        const f4 d = dropout_scale4(ph, (unsigned long long)(rowg + k0 + cq * 32 + i), a.stream_id, thr, ik);
        p[i] *= d.x; p[i + 1] *= d.y; p[i + 2] *= d.z; p[i + 3] *= d.w;
      }
    }
    store_slice_row<T>(sP, row, cq, p);
    fence_proxy_async_smem();
    tcgen05_fence_before();
    __syncthreads();
    if (t == 0) {
      mbar_wait(bar_v, j & 1);
      tcgen05_fence_after();
      const uint32_t pa = smem_u32(sP), va = smem_u32(sV);
#pragma unroll
      for (int s = 0; s < 2 * C::STEPS; ++s)
        tcgen05_mma_ss<C::KIND>(tmem + 128, desc_k<T>(pa, s), desc_mn<T>(va, 0, s), IDESC_O, s > 0 ? 1u : 0u);
      tcgen05_commit(bar_pv);
    }
    mbar_wait(bar_pv, j & 1);
    tcgen05_fence_after();
    {
      float v[16];
      ld16(tO, v);
#pragma unroll
      for (int i = 0; i < 16; ++i) o_acc[i] = o_acc[i] * alpha + v[i];
    }
    tcgen05_fence_before();
    __syncthreads();           // sRed, sMask, sK/sV/sP and TMEM are free for the next block
  }

  sRed[cq * 128 + row] = l_part;
  __syncthreads();
  const float l_run = sRed[row] + sRed[128 + row] + sRed[256 + row] + sRed[384 + row];
  {
    const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) o_acc[i] *= inv;
    stage_row16<T>(sP, row, cq, o_acc);                    // sP is idle: the last P V product has completed
    if (qlive && a.lse && cq == 0) a.lse[((long long)b * a.heads + h) * a.L + qrow] = m_run + __logf(fmaxf(l_run, 1e-37f));
  }
  fence_proxy_async_smem();
  tcgen05_fence_before();
  __syncthreads();
  if (t == 0) {
    store_tile<T>(&tmOut, sP, q0, h, b);                   // rows past L are clipped by the tensor map
    tma_store_commit();
    tma_store_wait_read<0>();
  }
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<256>(tmem);
  }
}

// =====================================================================================================
// backward for L <= 128, 512 threads, ONE kernel: dQ, dK, dV of one (batch, head).  grid = (heads, B)
// =====================================================================================================
// Stage 1 (tensor core): S = Q K^T, dP = dO V^T.
// Stage 2 (all threads): P = exp(S*scale + mask - lse), Pd = P∘drop, delta = rowsum(Pd∘dP) (== rowsum(dO∘O), so the
//                        forward output is not read), dS = P∘(dP∘drop - delta).
// Stage 3 (tensor core): dV = Pd^T dO, dQ = dS K, dK = dS^T Q; gradients leave through smem staging + TMA stores.
//   bf16: the row-major [q][k] images of Pd / dS serve both as K-major A (dQ) and, read MN-major, as the transposed
//         A operand (dV, dK); Q, K, dO tiles double as MN-major B operands.  smem = 4*16 + 2*32 = 128 KB.
//   fp32: MN-major fp32 operands need the 32-byte-atom swizzle, i.e. a different image.  Q/K/dO arrive a second time
//         through the atom-32 tensor maps (region R2, own barrier: they land while stages 1-2 run), and the transposed
//         A operands are written transposed by the threads into the region of the dead stage-1 tiles (R1), two images
//         at a time: {Pd^T, dS} first, then dS^T once dV has finished.  smem = 128 + 96 KB.
// TMEM: S [0,128)  dP [128,256)  dV [256,320)  dQ [320,384)  dK [384,448)
template <typename T>
__global__ void __launch_bounds__(kWideThreads, 1)
attn_bwd_block_kernel(const __grid_constant__ CUtensorMap tmQK, const __grid_constant__ CUtensorMap tmMN,
                      const __grid_constant__ CUtensorMap tmdO, const __grid_constant__ CUtensorMap tmdOmn,
                      const __grid_constant__ CUtensorMap tmOut, const AttnArgs a) {
  using C = AttCfg<T>;
  constexpr bool F32 = sizeof(T) == 4;
  constexpr int OPB = C::ATOMS * kTile;
  constexpr int NT = F32 ? 7 : 8;                                  // tiles of OPB bytes in the data region
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  if (F32 && smem != smem_raw) __trap();                           // fp32 has no slack to realign (see blk_smem)
  uint8_t* sQ = smem;
  uint8_t* sK = smem + OPB;
  uint8_t* sV = smem + 2 * OPB;
  uint8_t* sdO = smem + 3 * OPB;
  uint8_t* sKmn = F32 ? smem + 4 * OPB : sK;
  uint8_t* sQmn = F32 ? smem + 5 * OPB : sQ;
  uint8_t* sdOmn = F32 ? smem + 6 * OPB : sdO;
  uint8_t* sX = F32 ? smem : smem + 4 * OPB;                      // Pd (bf16) / Pd^T then dS^T (fp32)
  uint8_t* sY = F32 ? smem + 2 * OPB : smem + 6 * OPB;            // dS; first 2 KB double as the delta exchange
  float* sRed = reinterpret_cast<float*>(sY);                      // [4][128]
  float* sMask = reinterpret_cast<float*>(smem + NT * OPB);
  float* sLse = sMask + 128;
  uint64_t* bars = reinterpret_cast<uint64_t*>(sLse + 128);
  uint64_t *bar_qk = bars, *bar_dv = bars + 1, *bar_mn = bars + 2, *bar_s = bars + 3, *bar_a = bars + 4, *bar_b = bars + 5;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 6);

  const int t = threadIdx.x, warp = t >> 5, lane = t & 31;
  const int lq = warp & 3, cq = warp >> 2;
  const int row = lq * 32 + lane;
  const int h = blockIdx.x, b = blockIdx.y;
  const long long bh = (long long)b * a.heads + h;
  pdl_trigger();

  if (t == 0) {
    tma_prefetch_desc(&tmQK); tma_prefetch_desc(&tmMN); tma_prefetch_desc(&tmdO); tma_prefetch_desc(&tmdOmn);
    tma_prefetch_desc(&tmOut);
    mbar_init(bar_qk, 1); mbar_init(bar_dv, 1); mbar_init(bar_mn, 1);
    mbar_init(bar_s, 1); mbar_init(bar_a, 1); mbar_init(bar_b, 1);
    fence_barrier_init();
    pdl_wait();                                   // qkv / dO come from the previous kernels
    mbar_arrive_expect_tx(bar_qk, 2 * OPB);
    load_tile<T>(sQ, &tmQK, bar_qk, 0, h, b);
    load_tile<T>(sK, &tmQK, bar_qk, 0, a.heads + h, b);
    mbar_arrive_expect_tx(bar_dv, 2 * OPB);
    load_tile<T>(sdO, &tmdO, bar_dv, 0, h, b);
    load_tile<T>(sV, &tmQK, bar_dv, 0, 2 * a.heads + h, b);
    if constexpr (F32) {
      mbar_arrive_expect_tx(bar_mn, 3 * OPB);
      load_tile<T>(sdOmn, &tmdOmn, bar_mn, 0, h, b);
      load_tile<T>(sKmn, &tmMN, bar_mn, 0, a.heads + h, b);
      load_tile<T>(sQmn, &tmMN, bar_mn, 0, h, b);
    }
  }
  if (warp == 1) tmem_alloc<512>(tmem_slot);
  pdl_wait();
  if (t >= 128 && t < 256) {       // mask of key i, lse of query i
    const int i = t - 128;
    float mv = -3.0e38f, lse = 3.0e38f;
    if (i < a.L) {
      mv = a.mask ? (1.f - (float)a.mask[(long long)b * a.L + i]) * -10000.f : 0.f;
      lse = a.lse[bh * a.L + i];
    }
    sMask[i] = mv; sLse[i] = lse;
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem = *tmem_slot;
  const uint32_t tL = tmem + ((uint32_t)(lq * 32) << 16);

  constexpr uint32_t IDESC_S = make_idesc(C::FMT, false, false, 128, 128);
  constexpr uint32_t IDESC_KN = make_idesc(C::FMT, false, true, 128, 64);     // A K-major, B MN-major
  constexpr uint32_t IDESC_NN = make_idesc(C::FMT, true, true, 128, 64);      // A MN-major, B MN-major

  if (t == 0) {
    const uint32_t qa = smem_u32(sQ), ka = smem_u32(sK), va = smem_u32(sV), da = smem_u32(sdO);
    mbar_wait(bar_qk, 0);
    tcgen05_fence_after();
#pragma unroll
    for (int s = 0; s < C::STEPS; ++s)
      tcgen05_mma_ss<C::KIND>(tmem, desc_k<T>(qa, s), desc_k<T>(ka, s), IDESC_S, s > 0 ? 1u : 0u);
    mbar_wait(bar_dv, 0);
    tcgen05_fence_after();
#pragma unroll
    for (int s = 0; s < C::STEPS; ++s)
      tcgen05_mma_ss<C::KIND>(tmem + 128, desc_k<T>(da, s), desc_k<T>(va, s), IDESC_S, s > 0 ? 1u : 0u);
    tcgen05_commit(bar_s);
  }
  mbar_wait(bar_s, 0);
  tcgen05_fence_after();

  // ---- stage 2 --------------------------------------------------------------------------------------------------
  float p[32], dp[32];
  uint32_t keep = 0xffffffffu;
  ld32(tL + cq * 32, p);             // scores
  ld32(tL + 128 + cq * 32, dp);      // dP
  const float ik = a.p_drop > 0.f ? 1.f / (1.f - a.p_drop) : 1.f;
  float part = 0.f;
  {
    const float lse = sLse[row];
    if (a.p_drop > 0.f) {
      const Philox ph(*a.seed_dev);
      const uint32_t thr = dropout_threshold(a.p_drop);
      const unsigned long long idx = (unsigned long long)((bh * a.L + row) * a.Lp + cq * 32) >> 2;
      keep = 0u;
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        const uint4 r = ph(idx + i, a.stream_id);
        keep |= (r.x >= thr ? 1u : 0u) << (4 * i) | (r.y >= thr ? 2u : 0u) << (4 * i) | (r.z >= thr ? 4u : 0u) << (4 * i) |
                (r.w >= thr ? 8u : 0u) << (4 * i);
      }
    }
#pragma unroll
    for (int i = 0; i < 32; ++i) {
      const float mv = sMask[cq * 32 + i];
      p[i] = mv > -1.0e38f ? __expf(p[i] * a.scale + mv - lse) : 0.f;
      dp[i] = (keep >> i & 1u) ? dp[i] * ik : 0.f;               // dP through the dropout mask
      part += p[i] * dp[i];
    }
  }
  sRed[cq * 128 + row] = part;
  __syncthreads();
  const float delta = sRed[row] + sRed[128 + row] + sRed[256 + row] + sRed[384 + row];
  auto pd_at = [&](int i) { return (keep >> i & 1u) ? p[i] * ik : 0.f; };
  auto ds_at = [&](int i) { return p[i] * (dp[i] - delta); };
  if constexpr (F32) store_slice_transposed_gen(sX, row, cq, pd_at);
  else store_slice_row_gen<T>(sX, row, cq, pd_at);
  __syncthreads();                   // every thread has read sRed before dS lands on it
  store_slice_row_gen<T>(sY, row, cq, ds_at);
  fence_proxy_async_smem();
  tcgen05_fence_before();
  __syncthreads();

  // ---- stage 3 --------------------------------------------------------------------------------------------------
  if (t == 0) {
    tcgen05_fence_after();
    const uint32_t xa = smem_u32(sX), ya = smem_u32(sY);
    const uint32_t kb = smem_u32(sKmn), qb = smem_u32(sQmn), db = smem_u32(sdOmn);
    if constexpr (F32) {
      mbar_wait(bar_mn, 0);
      tcgen05_fence_after();
#pragma unroll
      for (int s = 0; s < 2 * C::STEPS; ++s)       // dV = (Pd^T) dO, A image already transposed
        tcgen05_mma_ss<C::KIND>(tmem + 256, desc_k<T>(xa, s), desc_mn<T>(db, 0, s), IDESC_KN, s > 0 ? 1u : 0u);
      tcgen05_commit(bar_a);                       // frees sX for dS^T as early as possible
#pragma unroll
      for (int s = 0; s < 2 * C::STEPS; ++s)       // dQ = dS K
        tcgen05_mma_ss<C::KIND>(tmem + 320, desc_k<T>(ya, s), desc_mn<T>(kb, 0, s), IDESC_KN, s > 0 ? 1u : 0u);
    } else {
#pragma unroll
      for (int s = 0; s < 2 * C::STEPS; ++s)       // dV = Pd^T dO, A read MN-major
        tcgen05_mma_ss<C::KIND>(tmem + 256, desc_mn<T>(xa, 0, s), desc_mn<T>(db, 0, s), IDESC_NN, s > 0 ? 1u : 0u);
#pragma unroll
      for (int s = 0; s < 2 * C::STEPS; ++s)       // dQ = dS K
        tcgen05_mma_ss<C::KIND>(tmem + 320, desc_k<T>(ya, s), desc_mn<T>(kb, 0, s), IDESC_KN, s > 0 ? 1u : 0u);
#pragma unroll
      for (int s = 0; s < 2 * C::STEPS; ++s)       // dK = dS^T Q, A read MN-major
        tcgen05_mma_ss<C::KIND>(tmem + 384, desc_mn<T>(ya, 0, s), desc_mn<T>(qb, 0, s), IDESC_NN, s > 0 ? 1u : 0u);
      tcgen05_commit(bar_a);
    }
  }
  mbar_wait(bar_a, 0);
  tcgen05_fence_after();
  if constexpr (F32) {
    store_slice_transposed_gen(sX, row, cq, ds_at);     // Pd^T is consumed: its space takes dS^T
    fence_proxy_async_smem();
    tcgen05_fence_before();
    __syncthreads();
    if (t == 0) {
      tcgen05_fence_after();
      const uint32_t xa = smem_u32(sX), qb = smem_u32(sQmn);
#pragma unroll
      for (int s = 0; s < 2 * C::STEPS; ++s)
        tcgen05_mma_ss<C::KIND>(tmem + 384, desc_k<T>(xa, s), desc_mn<T>(qb, 0, s), IDESC_KN, s > 0 ? 1u : 0u);
      tcgen05_commit(bar_b);
    }
    mbar_wait(bar_b, 0);
    tcgen05_fence_after();
  }

  // ---- epilogue: TMEM -> staging tiles (all operand tiles are dead now) -> TMA stores ---------------------------------
  {
    float g[16];
    ld16(tL + 320 + cq * 16, g);
#pragma unroll
    for (int i = 0; i < 16; ++i) g[i] *= a.scale;
    stage_row16<T>(smem, row, cq, g);                      // dQ
    ld16(tL + 384 + cq * 16, g);
#pragma unroll
    for (int i = 0; i < 16; ++i) g[i] *= a.scale;
    stage_row16<T>(smem + OPB, row, cq, g);                // dK
    ld16(tL + 256 + cq * 16, g);
    stage_row16<T>(smem + 2 * OPB, row, cq, g);            // dV
  }
  fence_proxy_async_smem();
  tcgen05_fence_before();
  __syncthreads();
  if (t == 0) {
    store_tile<T>(&tmOut, smem, 0, h, b);
    store_tile<T>(&tmOut, smem + OPB, 0, a.heads + h, b);
    store_tile<T>(&tmOut, smem + 2 * OPB, 0, 2 * a.heads + h, b);
    tma_store_commit();
    tma_store_wait_read<0>();
  }
  if (warp == 1) {
    tcgen05_fence_after();
    tmem_dealloc<512>(tmem);
  }
}

}  // namespace b200
