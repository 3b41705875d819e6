// gemm_kernel.cuh — the tcgen05 GEMM kernel, its launch variants and the per-type dispatch template.
// Included by gemm_sm100.cu (host logic, declares the three dispatch instantiations extern) and by the three
// gemm_inst_*.cu files, each of which instantiates the kernels of ONE operand/output type pair so that the
// instantiations compile in parallel.
#pragma once
#include "common.h"
#include "sm100_ptx.cuh"
#include "vec_utils.cuh"
#include "../include/b200seq.h"

#include <mutex>
#include <unordered_map>
#include <string>
#include <cstring>
#include <cstdlib>

namespace b200 {

// ------------------------------------------------------------------------------------------------
// device side
// ------------------------------------------------------------------------------------------------
struct GemmArgs {
  int M, N, K;
  int num_m, num_n, splits, kb_total, kb_per_split;
  int nb0, nbatch;   // batch = b1 * nb0 + b0 (e.g. b0 = head, b1 = sequence)
  const void* bias;
  const void* aux;
  long long ldaux;
  int act, aux_mode, accumulate, has_d2;
  float alpha;
  // attention-softmax epilogues (single N tile): 1 = P = softmax(alpha*acc + mask) [+ dropout copy to D2],
  // 2 = dS = P * (dPe - rowsum(P*dPe)) with P read as aux and dPe = acc through the regenerated dropout mask
  int epi_mode;
  const long long* attn_mask;
  float p_drop;
  const unsigned long long* seed_dev;
  unsigned long long stream_id;
  long long ldp;                 // row pitch of the probability tensors (dropout element index)
};

constexpr int kBM = 128;
constexpr int kGemmThreads = 192;        // 2 + 4 warps
constexpr int kStageBufBytes = 4096;  // 32 rows x 128 B
constexpr int kEpiBytes = 4 * 2 * kStageBufBytes;
constexpr int epi_bytes(int ew) { return ew * 2 * kStageBufBytes; }

// erf-GELU and its derivative from ONE exponential: with z = x / sqrt(2), exp(-z^2) = exp(-x^2 / 2) serves both the
// Abramowitz-Stegun 7.1.26 rational form of erf(z) (|error| <= 1.5e-7, far below TF32 / bf16 rounding) and the normal
// density.  The epilogue warps are issue-bound (one warp per scheduler), so instructions per element are what counts:
// erff() + __expf() cost roughly twice this.
__device__ __forceinline__ void gelu_pair(float x, float& g, float& dg) {
  const float e = __expf(-0.5f * x * x);
  const float az = fabsf(x) * 0.70710678118654752f;
  const float t = __fdividef(1.0f, fmaf(0.3275911f, az, 1.0f));
  const float poly = t * fmaf(t, fmaf(t, fmaf(t, fmaf(t, 1.061405429f, -1.453152027f), 1.421413741f), -0.284496736f), 0.254829592f);
  const float erf_abs = fmaf(-poly, e, 1.0f);                    // erf(|z|)
  const float cdf = 0.5f * (1.0f + copysignf(erf_abs, x));
  g = x * cdf;
  dg = fmaf(x * 0.3989422804014327f, e, cdf);
}
__device__ __forceinline__ float gelu_f(float x) { float g, dg; gelu_pair(x, g, dg); return g; }
__device__ __forceinline__ float dgelu_f(float x) { float g, dg; gelu_pair(x, g, dg); return dg; }

template <typename T>
__device__ __forceinline__ float to_f(T v);
template <>
__device__ __forceinline__ float to_f<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

// CTA2: the kernel is launched in clusters of two CTAs that share every tcgen05.mma (cta_group::2, M = 256):
// CTA r of the pair owns rows m0 + 128 r of the 256 x BN tile and loads only rows r*BN/2 .. of the B tile; the leader
// (r = 0) owns the smem-full / TMEM-empty barriers and issues the MMAs, commits are multicast to both CTAs.
// EW: number of epilogue warps, 4 or 8.  With 8, warps w and w+4 share a TMEM lane quadrant and take alternate column
// chunks of the tile; every warp keeps its own staging buffers, store ring and aux barriers.  The four-warp epilogue
// is issue-latency bound as soon as it does more than load-and-store (bias +6 us, GELU +14..25 us, a second output
// +10 us on the [4096 x 3072 x 768] GEMM, tools/gpu_probe_epilogue.py); two warps per scheduler hide each other's
// TMEM-load, MUFU and store-drain latencies.
template <typename TIn, typename TOut, int BN, bool A_MN, bool B_MN, int STAGES, bool EPI_SOFTMAX, bool CTA2 = false, int EW = 4>
__global__ void __launch_bounds__(64 + 32 * EW, 1)
gemm_tcgen05_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                    const __grid_constant__ CUtensorMap tmD, const __grid_constant__ CUtensorMap tmD2,
                    const __grid_constant__ CUtensorMap tmAux, const GemmArgs args) {
  constexpr int ES = sizeof(TIn);
  constexpr int BK = 128 / ES;        // elements of K per stage (one 128 B swizzle span)
  constexpr int UK = 32 / ES;         // K per tcgen05.mma (8 tf32 / 16 bf16)
  constexpr int EA = 128 / ES;        // elements per 128 B along MN (MN-major slab width)
  constexpr int A_BYTES = kBM * 128;
  constexpr int B_ROWS = CTA2 ? BN / 2 : BN;   // rows of the B tile held by this CTA
  constexpr int B_BYTES = B_ROWS * 128;
  static_assert(!(CTA2 && EPI_SOFTMAX), "softmax epilogues run single-CTA");
  static_assert(EW == 4 || (EW == 8 && !EPI_SOFTMAX), "4 or 8 epilogue warps; softmax epilogues own whole rows");
  constexpr int kEpiB = epi_bytes(EW);
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr int SLAB_BYTES = BK * 128;
  constexpr uint32_t ACC_STRIDE = (BN <= 64) ? 64 : (BN <= 128 ? 128 : 256);
  constexpr uint32_t TMEM_COLS = 2 * ACC_STRIDE;
  constexpr MmaKind KIND = (ES == 4) ? MmaKind::TF32 : MmaKind::F16;
  constexpr uint32_t IDESC = make_idesc(ES == 4 ? 2u : 1u, A_MN, B_MN, CTA2 ? 2 * kBM : kBM, BN);
  constexpr int OS = sizeof(TOut);
  constexpr int CW = 128 / OS;        // output columns per 128 B staging row (32 fp32 / 64 bf16)
  static_assert(BN % CW == 0, "BN must be a multiple of the staging chunk width");

  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
  uint8_t* smem_stage = smem;                                  // STAGES * STAGE_BYTES
  uint8_t* smem_epi = smem + STAGES * STAGE_BYTES;             // 4 warps x 2 x 4 KB
  uint64_t* bars = reinterpret_cast<uint64_t*>(smem_epi + kEpiB);
  uint64_t* full_bar = bars;
  uint64_t* empty_bar = bars + STAGES;
  uint64_t* tmem_full = bars + 2 * STAGES;
  uint64_t* tmem_empty = bars + 2 * STAGES + 2;
  uint64_t* aux_bar = bars + 2 * STAGES + 4;                   // [EW epilogue warps][2 staging buffers]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4 + 2 * EW);

  const int warp = threadIdx.x >> 5;
  const int lane = threadIdx.x & 31;
  const uint32_t crank = CTA2 ? cluster_ctarank() : 0u;      // 0 = leader of the pair
  pdl_trigger();                       // the next kernel on the stream may start its own prologue

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tmA);
    tma_prefetch_desc(&tmB);
    tma_prefetch_desc(&tmD);
    for (int s = 0; s < STAGES; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tmem_full[a], 1);
      mbar_init(&tmem_empty[a], CTA2 ? 2 * EW : EW);    // epilogue warps of both CTAs release the leader's slot
    }
    for (int a = 0; a < 2 * EW; ++a) mbar_init(&aux_bar[a], 1);
    fence_barrier_init();
  }
  if constexpr (CTA2) {
    cluster_sync_all();                  // both CTAs' barriers are initialised before anything can signal them
    if (warp == 1) tmem_alloc_2sm<TMEM_COLS>(tmem_slot);
  } else {
    if (warp == 1) tmem_alloc<TMEM_COLS>(tmem_slot);
  }
  tcgen05_fence_before();
  __syncthreads();
  tcgen05_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  pdl_wait();                          // everything above overlapped the previous kernel's tail; its results from here on

  // CTA2: args.num_m counts 256-row pair tiles and the pair index replaces blockIdx.x in the tile schedule
  const int num_tiles = args.nbatch * args.num_m * args.num_n * args.splits;
  const int sched_id = CTA2 ? (int)(blockIdx.x >> 1) : (int)blockIdx.x;
  const int sched_n = CTA2 ? (int)(gridDim.x >> 1) : (int)gridDim.x;
  constexpr int TILE_M = CTA2 ? 2 * kBM : kBM;

  if (warp == 0) {
    // ===================================== TMA producer =====================================
    if (lane == 0) {
      uint32_t it = 0;
      for (int tile = sched_id; tile < num_tiles; tile += sched_n) {
        const int split = tile % args.splits;
        int rest = tile / args.splits;
        const int n0 = (rest % args.num_n) * BN + (int)crank * B_ROWS;          // this CTA's share of the B tile
        rest /= args.num_n;
        const int m0 = (rest % args.num_m) * TILE_M + (int)crank * kBM;       // this CTA's 128 rows
        const int bt = rest / args.num_m;
        const int b0 = bt % args.nb0, b1 = bt / args.nb0;
        const int kb0 = split * args.kb_per_split;
        const int kb1 = min(args.kb_total, kb0 + args.kb_per_split);
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const uint32_t s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&empty_bar[s], ph ^ 1);
          uint8_t* sA = smem_stage + s * STAGE_BYTES;
          uint8_t* sB = sA + A_BYTES;
          const int k0 = kb * BK;
          if constexpr (CTA2) {
            // the leader's barrier collects the bytes of both CTAs; only the leader arms it
            if (crank == 0) mbar_arrive_expect_tx(&full_bar[s], 2 * STAGE_BYTES);
            const uint32_t fb = mapa_u32(smem_u32(&full_bar[s]), 0);
            if constexpr (!A_MN) {
              tma_load_4d_2sm(sA, &tmA, fb, k0, m0, b0, b1);
            } else {
#pragma unroll
              for (int sl = 0; sl < kBM / EA; ++sl) tma_load_4d_2sm(sA + sl * SLAB_BYTES, &tmA, fb, m0 + sl * EA, k0, b0, b1);
            }
            if constexpr (!B_MN) {
              tma_load_4d_2sm(sB, &tmB, fb, k0, n0, b0, b1);
            } else {
#pragma unroll
              for (int sl = 0; sl < B_ROWS / EA; ++sl) tma_load_4d_2sm(sB + sl * SLAB_BYTES, &tmB, fb, n0 + sl * EA, k0, b0, b1);
            }
          } else {
            mbar_arrive_expect_tx(&full_bar[s], STAGE_BYTES);
            if constexpr (!A_MN) {
              tma_load_4d(sA, &tmA, &full_bar[s], k0, m0, b0, b1);
            } else {
#pragma unroll
              for (int sl = 0; sl < kBM / EA; ++sl) tma_load_4d(sA + sl * SLAB_BYTES, &tmA, &full_bar[s], m0 + sl * EA, k0, b0, b1);
            }
            if constexpr (!B_MN) {
              tma_load_4d(sB, &tmB, &full_bar[s], k0, n0, b0, b1);
            } else {
#pragma unroll
              for (int sl = 0; sl < BN / EA; ++sl) tma_load_4d(sB + sl * SLAB_BYTES, &tmB, &full_bar[s], n0 + sl * EA, k0, b0, b1);
            }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ===================================== MMA issuer =======================================
    if (lane == 0 && crank == 0) {
      uint32_t it = 0;
      uint32_t tcount = 0;
      for (int tile = sched_id; tile < num_tiles; tile += sched_n, ++tcount) {
        const int split = tile % args.splits;
        const int kb0 = split * args.kb_per_split;
        const int kb1 = min(args.kb_total, kb0 + args.kb_per_split);
        const uint32_t acc = tcount & 1;
        mbar_wait(&tmem_empty[acc], ((tcount >> 1) & 1) ^ 1);
        tcgen05_fence_after();
        const uint32_t d_tmem = tmem_base + acc * ACC_STRIDE;
        for (int kb = kb0; kb < kb1; ++kb, ++it) {
          const uint32_t s = it % STAGES;
          const uint32_t ph = (it / STAGES) & 1;
          mbar_wait(&full_bar[s], ph);
          tcgen05_fence_after();
          const uint32_t a_addr = smem_u32(smem_stage + s * STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_BYTES;
          // MN-major fp32 tiles use the 32-byte-atom swizzle (layout type 1, 4-row atoms)
          constexpr uint32_t MN_SBO = (ES == 4) ? 512 : 1024;
          constexpr uint32_t MN_LT = (ES == 4) ? 1 : 2;
          const uint64_t a_desc = A_MN ? make_smem_desc_sw128(a_addr, SLAB_BYTES, MN_SBO, MN_LT)
                                       : make_smem_desc_sw128(a_addr, 16, 1024);
          const uint64_t b_desc = B_MN ? make_smem_desc_sw128(b_addr, SLAB_BYTES, MN_SBO, MN_LT)
                                       : make_smem_desc_sw128(b_addr, 16, 1024);
#pragma unroll
          for (int j = 0; j < BK / UK; ++j) {
            // advance inside the stage: K-major -> 32 B along the swizzled row; MN-major -> UK rows of 128 B
            const uint64_t a_off = A_MN ? (uint64_t)((j * UK * 128) >> 4) : (uint64_t)((j * 32) >> 4);
            const uint64_t b_off = B_MN ? (uint64_t)((j * UK * 128) >> 4) : (uint64_t)((j * 32) >> 4);
            if constexpr (CTA2) tcgen05_mma_ss_2sm<KIND>(d_tmem, a_desc + a_off, b_desc + b_off, IDESC, (kb > kb0 || j > 0) ? 1u : 0u);
            else tcgen05_mma_ss<KIND>(d_tmem, a_desc + a_off, b_desc + b_off, IDESC, (kb > kb0 || j > 0) ? 1u : 0u);
          }
          // frees the smem stage (in both CTAs of a pair) once these MMAs have read it
          if constexpr (CTA2) tcgen05_commit_2sm(&empty_bar[s], 3);
          else tcgen05_commit(&empty_bar[s]);
        }
        // accumulator complete -> epilogue (of both CTAs)
        if constexpr (CTA2) tcgen05_commit_2sm(&tmem_full[acc], 3);
        else tcgen05_commit(&tmem_full[acc]);
      }
    }
  } else {
    // ===================================== epilogue =========================================
    const int ew = warp - 2;               // staging-buffer owner index 0..EW-1
    constexpr int CSTEP = EW / 4;          // column chunks are dealt round-robin to the warps of a quadrant
    const int c_first = ew >> 2;
    const int q = warp & 3;                // TMEM lane quarter this warp may read
    uint8_t* stg = smem_epi + ew * 2 * kStageBufBytes;
    uint32_t ev = 0;                       // store-event counter (ring of 2 staging buffers)
    uint32_t tcount = 0;
    const TOut* bias = reinterpret_cast<const TOut*>(args.bias);
    const TOut* aux = reinterpret_cast<const TOut*>(args.aux);
    const bool vec_ok = (args.N % (16 / OS)) == 0;
    // aux tiles (residual / pre-activation) are fetched by TMA into the staging buffer the result will leave
    // from, one chunk ahead of the math; the row-owner then reads its 128 B row from swizzled smem.
    const bool tma_aux = (args.aux_mode != B200_AUX_NONE) && !args.has_d2;
    uint64_t* my_aux_bar = aux_bar + ew * 2;
    int cur_b0 = 0, cur_b1 = 0;
    auto issue_aux = [&](int nc_, int m0_, uint32_t ev_) {       // lane 0 only
      tma_store_wait_read<0>();                                  // every earlier store has left its buffer
      uint64_t* bar = &my_aux_bar[ev_ & 1];
      mbar_arrive_expect_tx(bar, kStageBufBytes);
      tma_load_4d(stg + (ev_ & 1) * kStageBufBytes, &tmAux, bar, nc_, m0_, cur_b0, cur_b1);
    };
    uint32_t ax = 0;                       // aux-tile sequence counter of the softmax-backward epilogue

    for (int tile = sched_id; tile < num_tiles; tile += sched_n, ++tcount) {
      const int split = tile % args.splits;
      int rest = tile / args.splits;
      const int n0 = (rest % args.num_n) * BN;
      rest /= args.num_n;
      const int m0 = (rest % args.num_m) * TILE_M + (int)crank * kBM;
      const int bt = rest / args.num_m;
      const int b0 = bt % args.nb0, b1 = bt / args.nb0;
      const uint32_t acc = tcount & 1;
      mbar_wait(&tmem_full[acc], (tcount >> 1) & 1);
      tcgen05_fence_after();
      const int row = m0 + q * 32 + lane;
      const bool add_bias = (bias != nullptr) && (split == 0);
      cur_b0 = b0; cur_b1 = b1;
      if constexpr (EPI_SOFTMAX) {
        // ======================= attention softmax epilogues (whole row inside this tile) =======================
        const bool live = (m0 + q * 32 < args.M);
        const int NCH = (args.N + CW - 1) / CW;
        const long long rowg = ((long long)bt * args.M + row) * args.ldp;      // element index of (row, 0)
        const Philox ph(args.seed_dev ? *args.seed_dev : 0ull);
        const uint32_t thr = dropout_threshold(args.p_drop);
        const float ik = args.p_drop > 0.f ? 1.f / (1.f - args.p_drop) : 1.f;
        const uint32_t tbase = tmem_base + ((uint32_t)(q * 32) << 16) + acc * ACC_STRIDE;
        auto load_chunk = [&](int c, float (&v)[CW]) {
#pragma unroll
          for (int h = 0; h < CW / 32; ++h) {
            uint32_t r[32];
            tmem_ld_32x32(tbase + c * CW + h * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) v[h * 32 + j] = __uint_as_float(r[j]);
          }
        };
        auto store_chunk = [&](const CUtensorMap* tm, const float (&v)[CW], int c, uint32_t bufsel) {
          uint8_t* buf = stg + (bufsel & 1) * kStageBufBytes;
          if (lane == 0) tma_store_wait_read<1>();
          __syncwarp();
          uint8_t* rowp = buf + lane * 128;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            uint4 u;
            if constexpr (OS == 4) {
              u.x = __float_as_uint(v[4 * i + 0]); u.y = __float_as_uint(v[4 * i + 1]);
              u.z = __float_as_uint(v[4 * i + 2]); u.w = __float_as_uint(v[4 * i + 3]);
            } else {
              __nv_bfloat162 p0 = __floats2bfloat162_rn(v[8 * i + 0], v[8 * i + 1]);
              __nv_bfloat162 p1 = __floats2bfloat162_rn(v[8 * i + 2], v[8 * i + 3]);
              __nv_bfloat162 p2 = __floats2bfloat162_rn(v[8 * i + 4], v[8 * i + 5]);
              __nv_bfloat162 p3 = __floats2bfloat162_rn(v[8 * i + 6], v[8 * i + 7]);
              u.x = *reinterpret_cast<uint32_t*>(&p0); u.y = *reinterpret_cast<uint32_t*>(&p1);
              u.z = *reinterpret_cast<uint32_t*>(&p2); u.w = *reinterpret_cast<uint32_t*>(&p3);
            }
            *reinterpret_cast<uint4*>(rowp + ((i ^ (lane & 7)) << 4)) = u;
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            tma_store_4d(tm, buf, n0 + c * CW, m0 + q * 32, b0, b1);
            tma_store_commit();
          }
        };
        if (live && args.epi_mode == 1) {
          // ---- forward: online (max, sum) over the row, then normalise, optionally write the dropped copy ----
          const long long* mrow = args.attn_mask ? args.attn_mask + (long long)b1 * args.N : nullptr;
          float m = -3.0e38f, l = 0.f;
#pragma unroll 1
          for (int c = 0; c < NCH; ++c) {
            float v[CW];
            load_chunk(c, v);
            float cm = -3.0e38f;
#pragma unroll
            for (int j = 0; j < CW; ++j) {
              const int col = c * CW + j;
              const float madd = (mrow && col < args.N) ? (1.f - (float)__ldg(mrow + col)) * -10000.f : 0.f;
              v[j] = col < args.N ? v[j] * args.alpha + madd : -3.0e38f;
              cm = fmaxf(cm, v[j]);
            }
            const float mn = fmaxf(m, cm);
            float sum = 0.f;
#pragma unroll
            for (int j = 0; j < CW; ++j) sum += (c * CW + j < args.N) ? __expf(v[j] - mn) : 0.f;
            l = l * __expf(m - mn) + sum;
            m = mn;
          }
          const float inv = 1.f / l;
#pragma unroll 1
          for (int c = 0; c < NCH; ++c) {
            float v[CW];
            load_chunk(c, v);
#pragma unroll
            for (int j = 0; j < CW; ++j) {
              const int col = c * CW + j;
              const float madd = (mrow && col < args.N) ? (1.f - (float)__ldg(mrow + col)) * -10000.f : 0.f;
              v[j] = col < args.N ? __expf(v[j] * args.alpha + madd - m) * inv : 0.f;
            }
            store_chunk(&tmD, v, c, ev++);
            if (args.has_d2) {
#pragma unroll
              for (int j = 0; j < CW; j += 4) {
                const f4 d = dropout_scale4(ph, (unsigned long long)(rowg + c * CW + j), args.stream_id, thr, ik);
                v[j] *= d.x; v[j + 1] *= d.y; v[j + 2] *= d.z; v[j + 3] *= d.w;
              }
              store_chunk(&tmD2, v, c, ev++);
            }
          }
        } else if (live) {
          // ---- backward: two passes over (acc = dP, aux = P); aux tiles stream through the staging buffers ----
          auto aux_issue = [&](int c, uint32_t seq) {           // lane 0 only
            tma_store_wait_read<0>();
            uint64_t* bar = &my_aux_bar[seq & 1];
            mbar_arrive_expect_tx(bar, kStageBufBytes);
            tma_load_4d(stg + (seq & 1) * kStageBufBytes, &tmAux, bar, n0 + c * CW, m0 + q * 32, b0, b1);
          };
          auto aux_read = [&](uint32_t seq, float (&pv)[CW]) {
            mbar_wait(&my_aux_bar[seq & 1], (seq >> 1) & 1);
            const uint8_t* rowp = stg + (seq & 1) * kStageBufBytes + lane * 128;
#pragma unroll
            for (int i = 0; i < 8; ++i) {
              const uint4 u = *reinterpret_cast<const uint4*>(rowp + ((i ^ (lane & 7)) << 4));
              const TOut* a = reinterpret_cast<const TOut*>(&u);
#pragma unroll
              for (int e = 0; e < 16 / OS; ++e) pv[i * (16 / OS) + e] = to_f<TOut>(a[e]);
            }
            __syncwarp();
          };
          if (lane == 0) aux_issue(0, ax);
          float dot = 0.f;
#pragma unroll 1
          for (int pass = 0; pass < 2; ++pass) {
#pragma unroll 1
            for (int c = 0; c < NCH; ++c) {
              float pv[CW], v[CW];
              aux_read(ax, pv);
              const bool more = !(pass == 1 && c == NCH - 1);
              if (lane == 0 && more) aux_issue(c + 1 < NCH ? c + 1 : 0, ax + 1);
              load_chunk(c, v);
              if (args.p_drop > 0.f) {
#pragma unroll
                for (int j = 0; j < CW; j += 4) {
                  const f4 d = dropout_scale4(ph, (unsigned long long)(rowg + c * CW + j), args.stream_id, thr, ik);
                  v[j] *= d.x; v[j + 1] *= d.y; v[j + 2] *= d.z; v[j + 3] *= d.w;
                }
              }
              if (pass == 0) {
#pragma unroll
                for (int j = 0; j < CW; ++j) dot += pv[j] * v[j];
              } else {
#pragma unroll
                for (int j = 0; j < CW; ++j) v[j] = pv[j] * (v[j] - dot);
                store_chunk(&tmD, v, c, ax);
              }
              ++ax;
            }
          }
        }
        tcgen05_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive(&tmem_empty[acc]);
        continue;
      }
      if (tma_aux && lane == 0 && n0 + c_first * CW < args.N && m0 + q * 32 < args.M) issue_aux(n0 + c_first * CW, m0 + q * 32, ev);

#pragma unroll 1
      for (int c = c_first; c < BN / CW; c += CSTEP) {
        const int nc = n0 + c * CW;
        if (nc >= args.N || m0 + q * 32 >= args.M) break;   // warp-uniform: nothing of this chunk is in range
        float v[CW];
#pragma unroll
        for (int h = 0; h < CW / 32; ++h) {
          uint32_t r[32];
          tmem_ld_32x32(tmem_base + ((uint32_t)(q * 32) << 16) + acc * ACC_STRIDE + c * CW + h * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) v[h * 32 + j] = __uint_as_float(r[j]) * args.alpha;
        }
        if (add_bias) {
          if (vec_ok) {
#pragma unroll
            for (int j = 0; j < CW; j += 16 / OS) {
              if (nc + j < args.N) {
                const uint4 u = __ldg(reinterpret_cast<const uint4*>(bias + nc + j));
                const TOut* b = reinterpret_cast<const TOut*>(&u);
#pragma unroll
                for (int e = 0; e < 16 / OS; ++e) v[j + e] += to_f<TOut>(b[e]);
              }
            }
          } else {
#pragma unroll
            for (int j = 0; j < CW; ++j)
              if (nc + j < args.N) v[j] += to_f<TOut>(__ldg(bias + nc + j));
          }
        }

        auto stage_and_store = [&](const CUtensorMap* tm, bool reduce, const float (&v)[CW]) {
          uint8_t* buf = stg + (ev & 1) * kStageBufBytes;
          if (lane == 0) tma_store_wait_read<1>();   // the store that last used this buffer has drained
          __syncwarp();
          uint8_t* rowp = buf + lane * 128;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            uint4 u;
            if constexpr (OS == 4) {
              u.x = __float_as_uint(v[4 * i + 0]);
              u.y = __float_as_uint(v[4 * i + 1]);
              u.z = __float_as_uint(v[4 * i + 2]);
              u.w = __float_as_uint(v[4 * i + 3]);
            } else {
              __nv_bfloat162 p0 = __floats2bfloat162_rn(v[8 * i + 0], v[8 * i + 1]);
              __nv_bfloat162 p1 = __floats2bfloat162_rn(v[8 * i + 2], v[8 * i + 3]);
              __nv_bfloat162 p2 = __floats2bfloat162_rn(v[8 * i + 4], v[8 * i + 5]);
              __nv_bfloat162 p3 = __floats2bfloat162_rn(v[8 * i + 6], v[8 * i + 7]);
              u.x = *reinterpret_cast<uint32_t*>(&p0);
              u.y = *reinterpret_cast<uint32_t*>(&p1);
              u.z = *reinterpret_cast<uint32_t*>(&p2);
              u.w = *reinterpret_cast<uint32_t*>(&p3);
            }
            *reinterpret_cast<uint4*>(rowp + ((i ^ (lane & 7)) << 4)) = u;
          }
          fence_proxy_async_smem();
          __syncwarp();
          if (lane == 0) {
            if (reduce) tma_reduce_add_4d(tm, buf, nc, m0 + q * 32, b0, b1);
            else tma_store_4d(tm, buf, nc, m0 + q * 32, b0, b1);
            tma_store_commit();
          }
          ++ev;
        };

        if (args.act == B200_ACT_GELU_SAVE_GRAD) {
          float dv[CW];                                  // D2 = gelu'(x) for the backward pass, D = gelu(x)
#pragma unroll
          for (int j = 0; j < CW; ++j) gelu_pair(v[j], v[j], dv[j]);
          stage_and_store(&tmD2, false, dv);
        } else if (args.has_d2) {
          stage_and_store(&tmD2, false, v);
        }

        if (args.act == B200_ACT_GELU) {
#pragma unroll
          for (int j = 0; j < CW; ++j) v[j] = gelu_f(v[j]);
        } else if (args.act == B200_ACT_TANH) {
#pragma unroll
          for (int j = 0; j < CW; ++j) v[j] = tanhf(v[j]);
        }

        if (tma_aux) {
          mbar_wait(&my_aux_bar[ev & 1], (ev >> 1) & 1);
          const uint8_t* rowp = stg + (ev & 1) * kStageBufBytes + lane * 128;
#pragma unroll
          for (int i = 0; i < 8; ++i) {
            const uint4 u = *reinterpret_cast<const uint4*>(rowp + ((i ^ (lane & 7)) << 4));
            const TOut* a = reinterpret_cast<const TOut*>(&u);
#pragma unroll
            for (int e = 0; e < 16 / OS; ++e) {
              const float x = to_f<TOut>(a[e]);
              const int j = i * (16 / OS) + e;
              if (args.aux_mode == B200_AUX_ADD) v[j] += x;
              else if (args.aux_mode == B200_AUX_MUL) v[j] *= x;
              else if (args.aux_mode == B200_AUX_MUL_DGELU) v[j] *= dgelu_f(x);
              else v[j] *= (1.0f - x * x);
            }
          }
          __syncwarp();                                           // everyone has consumed the aux tile
          const int nc_next = nc + CSTEP * CW;
          if (lane == 0 && c + CSTEP < BN / CW && nc_next < args.N) issue_aux(nc_next, m0 + q * 32, ev + 1);
        } else if (args.aux_mode != B200_AUX_NONE && row < args.M) {
          const TOut* arow = aux + (long long)row * args.ldaux + nc;
#pragma unroll
          for (int j = 0; j < CW; j += 16 / OS) {
            if (nc + j < args.N) {
              const uint4 u = __ldg(reinterpret_cast<const uint4*>(arow + j));
              const TOut* a = reinterpret_cast<const TOut*>(&u);
#pragma unroll
              for (int e = 0; e < 16 / OS; ++e) {
                const float x = to_f<TOut>(a[e]);
                if (args.aux_mode == B200_AUX_ADD) v[j + e] += x;
                else if (args.aux_mode == B200_AUX_MUL) v[j + e] *= x;
                else if (args.aux_mode == B200_AUX_MUL_DGELU) v[j + e] *= dgelu_f(x);
                else v[j + e] *= (1.0f - x * x);
              }
            }
          }
        }
        stage_and_store(&tmD, args.accumulate != 0, v);
      }
      // all TMEM reads of this accumulator slot are done -> hand it back to the MMA warp
      tcgen05_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (CTA2) {
          if (crank == 0) mbar_arrive(&tmem_empty[acc]);
          else mbar_arrive_cluster(mapa_u32(smem_u32(&tmem_empty[acc]), 0));
        } else {
          mbar_arrive(&tmem_empty[acc]);
        }
      }
    }
    if (lane == 0) tma_store_wait_read<0>();
    __syncwarp();
  }

  tcgen05_fence_before();
  if constexpr (CTA2) {
    cluster_sync_all();       // neither CTA may free TMEM (or exit, its smem is read by the peer's MMAs) before both are done
    if (warp == 1) {
      tcgen05_fence_after();
      tmem_dealloc_2sm<TMEM_COLS>(tmem_base);
    }
  } else {
    __syncthreads();
    if (warp == 1) {
      tcgen05_fence_after();
      tmem_dealloc<TMEM_COLS>(tmem_base);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// host side: dispatch
// ------------------------------------------------------------------------------------------------
template <int BN, bool CTA2 = false, int EW = 4>
constexpr int stages_for() {
  // 227 KB usable; keep 8 KB per epilogue warp for its staging ring, ~2 KB for alignment slack + barriers.
  constexpr int stage = kBM * 128 + (CTA2 ? BN / 2 : BN) * 128;
  constexpr int avail = 232448 - epi_bytes(EW) - 2048;
  constexpr int s = avail / stage;
  return s > 8 ? 8 : s;
}

// One launch of a (CTA2, EW) variant.  CTA2: clusters of two CTAs, one pair per TPC, 256 x BN tiles.
template <typename TIn, typename TOut, int BN, bool A_MN, bool B_MN, bool CTA2, int EW>
static int launch_one(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tD, const CUtensorMap& tD2,
                      const CUtensorMap& tX, const GemmArgs& args, cudaStream_t stream) {
  constexpr int STAGES = stages_for<BN, CTA2, EW>();
  constexpr int smem = STAGES * (kBM * 128 + (CTA2 ? BN / 2 : BN) * 128) + epi_bytes(EW) + 1024 /*align*/ +
                       (2 * STAGES + 4 + 2 * EW) * 8 + 16;
  auto kern = gemm_tcgen05_kernel<TIn, TOut, BN, A_MN, B_MN, STAGES, false, CTA2, EW>;
  static bool attr_set = false;
  if (!attr_set) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set = true;
  }
  const long long tiles = (long long)args.nbatch * args.num_m * args.num_n * args.splits;     // pair tiles when CTA2
  const int units = CTA2 ? compute_sms() / 2 : compute_sms();
  const int grid = (CTA2 ? 2 : 1) * (int)(tiles < units ? tiles : units);
  B200_CUDA(launch_pdl(kern, dim3(grid), dim3(64 + 32 * EW), smem, stream, CTA2 ? 2 : 1, tA, tB, tD, tD2, tX, args));
  B200_LAUNCHED("gemm_tcgen05_kernel");
  return 0;
}

// Epilogues that do more than load-and-store get eight epilogue warps (see the kernel's EW comment); B200_GEMM_EW=4/8
// forces one variant.  Only instantiated for K-major A with matching in/out types (forward and data-gradient GEMMs).
template <typename TIn, typename TOut, bool A_MN>
static bool wide_epilogue(const GemmArgs& args) {
  if constexpr (A_MN || sizeof(TIn) != sizeof(TOut)) {
    return false;
  } else {
    const char* e = getenv("B200_GEMM_EW");
    if (e) return atoi(e) == 8;
    return args.bias != nullptr || args.act != B200_ACT_NONE || args.aux_mode != B200_AUX_NONE || args.has_d2;
  }
}

template <typename TIn, typename TOut, int BN, bool A_MN, bool B_MN>
static int launch_gemm_cta2(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tD, const CUtensorMap& tD2,
                            const CUtensorMap& tX, const GemmArgs& args, cudaStream_t stream) {
  if constexpr (!A_MN && sizeof(TIn) == sizeof(TOut)) {
    if (wide_epilogue<TIn, TOut, A_MN>(args)) return launch_one<TIn, TOut, BN, A_MN, B_MN, true, 8>(tA, tB, tD, tD2, tX, args, stream);
  }
  return launch_one<TIn, TOut, BN, A_MN, B_MN, true, 4>(tA, tB, tD, tD2, tX, args, stream);
}

template <typename TIn, typename TOut, int BN, bool A_MN, bool B_MN>
static int launch_gemm(const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tD, const CUtensorMap& tD2,
                       const CUtensorMap& tX, const GemmArgs& args, cudaStream_t stream) {
  if constexpr (!A_MN && !B_MN && sizeof(TIn) == sizeof(TOut)) {
    if (args.epi_mode != 0) {        // attention-softmax epilogue kernels are separate instantiations
      constexpr int STAGES = stages_for<BN>();
      constexpr int smem = STAGES * (kBM * 128 + BN * 128) + kEpiBytes + 1024 /*align*/ + (2 * STAGES + 12) * 8 + 16;
      auto kern_sm = gemm_tcgen05_kernel<TIn, TOut, BN, A_MN, B_MN, STAGES, true>;
      static bool attr_set_sm = false;
      if (!attr_set_sm) {
        B200_CUDA(cudaFuncSetAttribute(kern_sm, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
        attr_set_sm = true;
      }
      const long long tiles_sm = (long long)args.nbatch * args.num_m * args.num_n * args.splits;
      const int grid_sm = tiles_sm < num_sms() ? (int)tiles_sm : num_sms();
      B200_CUDA(launch_pdl(kern_sm, dim3(grid_sm), dim3(kGemmThreads), smem, stream, 1, tA, tB, tD, tD2, tX, args));
      B200_LAUNCHED("gemm_tcgen05_kernel<softmax>");
      return 0;
    }
  }
  if (args.epi_mode != 0) return fail(B200_ERR_UNSUPPORTED, "gemm: softmax epilogues need K-major operands and matching in/out dtype");
  if constexpr (!A_MN && sizeof(TIn) == sizeof(TOut)) {
    if (wide_epilogue<TIn, TOut, A_MN>(args)) return launch_one<TIn, TOut, BN, A_MN, B_MN, false, 8>(tA, tB, tD, tD2, tX, args, stream);
  }
  return launch_one<TIn, TOut, BN, A_MN, B_MN, false, 4>(tA, tB, tD, tD2, tX, args, stream);
}

template <typename TIn, typename TOut, int BN>
static int dispatch_major_cta2(bool a_mn, bool b_mn, const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tD,
                               const CUtensorMap& tD2, const CUtensorMap& tX, const GemmArgs& args, cudaStream_t stream) {
  if (!a_mn && !b_mn) return launch_gemm_cta2<TIn, TOut, BN, false, false>(tA, tB, tD, tD2, tX, args, stream);
  if (!a_mn && b_mn) return launch_gemm_cta2<TIn, TOut, BN, false, true>(tA, tB, tD, tD2, tX, args, stream);
  if (a_mn && b_mn) return launch_gemm_cta2<TIn, TOut, BN, true, true>(tA, tB, tD, tD2, tX, args, stream);
  return launch_gemm_cta2<TIn, TOut, BN, true, false>(tA, tB, tD, tD2, tX, args, stream);
}

template <typename TIn, typename TOut, int BN>
static int dispatch_major(bool a_mn, bool b_mn, const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tD,
                          const CUtensorMap& tD2, const CUtensorMap& tX, const GemmArgs& args, cudaStream_t stream) {
  if (!a_mn && !b_mn) return launch_gemm<TIn, TOut, BN, false, false>(tA, tB, tD, tD2, tX, args, stream);
  if (!a_mn && b_mn) return launch_gemm<TIn, TOut, BN, false, true>(tA, tB, tD, tD2, tX, args, stream);
  if (a_mn && b_mn) return launch_gemm<TIn, TOut, BN, true, true>(tA, tB, tD, tD2, tX, args, stream);
  return launch_gemm<TIn, TOut, BN, true, false>(tA, tB, tD, tD2, tX, args, stream);
}

template <typename TIn, typename TOut>
int dispatch_bn(int bn, bool cta2, bool a_mn, bool b_mn, const CUtensorMap& tA, const CUtensorMap& tB, const CUtensorMap& tD,
                       const CUtensorMap& tD2, const CUtensorMap& tX, const GemmArgs& args, cudaStream_t stream) {
  if (cta2) {
    switch (bn) {
      case 128: return dispatch_major_cta2<TIn, TOut, 128>(a_mn, b_mn, tA, tB, tD, tD2, tX, args, stream);
      case 256: return dispatch_major_cta2<TIn, TOut, 256>(a_mn, b_mn, tA, tB, tD, tD2, tX, args, stream);
    }
    return fail(B200_ERR_ARG, "cta-pair GEMM supports BN 128 / 256, not %d", bn);
  }
  switch (bn) {
    case 64: return dispatch_major<TIn, TOut, 64>(a_mn, b_mn, tA, tB, tD, tD2, tX, args, stream);
    case 128: return dispatch_major<TIn, TOut, 128>(a_mn, b_mn, tA, tB, tD, tD2, tX, args, stream);
    case 192: return dispatch_major<TIn, TOut, 192>(a_mn, b_mn, tA, tB, tD, tD2, tX, args, stream);
    case 256: return dispatch_major<TIn, TOut, 256>(a_mn, b_mn, tA, tB, tD, tD2, tX, args, stream);
  }
  return fail(B200_ERR_ARG, "unsupported BN %d", bn);
}


}  // namespace b200
