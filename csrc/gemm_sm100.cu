// gemm_sm100.cu — persistent, warp-specialised tcgen05 GEMM for sm_100a.
//
//   D[M,N] = epilogue( alpha * sum_k A(m,k) * B(n,k) )
//
// Design (B200-first, nothing here comes from the reference, which calls cuBLAS through nn.Linear:
// hetseq/bert_modeling.py:166-172,352-354,388,424,515,526,544):
//   * one CTA per SM, static round-robin tile scheduler over (m-block, n-block, k-split);
//   * warp 0  : TMA producer — fills a ring of {A,B} shared-memory stages (128-byte swizzle);
//   * warp 1  : allocates TMEM and issues tcgen05.mma (one elected lane), accumulating a 128xBN fp32
//               tile in one of two TMEM accumulator slots so the epilogue of tile i overlaps tile i+1;
//   * warps 2-5: epilogue — tcgen05.ld their 32-row quarter, apply alpha/bias/activation/aux in
//               registers, stage 32x128B through swizzled smem, and TMA-store (or TMA reduce-add).
//   * operands may be K-major or MN-major in global memory; MN-major tiles are fetched as 128-byte-wide
//     slabs and described to the tensor core with the MN-major canonical layout, so forward, dgrad and
//     wgrad all run through this kernel without any transpose pass.
//   * kind::tf32 consumes fp32 storage directly (fp32/TF32 parity mode); kind::f16 consumes bf16.
#include "common.h"
#include "gemm_kernel.cuh"

namespace b200 {

// instantiated in gemm_inst_{tf32,bf16,bf16f32}.cu
extern template int dispatch_bn<float, float>(int, bool, bool, bool, const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                              const CUtensorMap&, const CUtensorMap&, const GemmArgs&, cudaStream_t);
extern template int dispatch_bn<__nv_bfloat16, __nv_bfloat16>(int, bool, bool, bool, const CUtensorMap&, const CUtensorMap&,
                                                              const CUtensorMap&, const CUtensorMap&, const CUtensorMap&,
                                                              const GemmArgs&, cudaStream_t);
extern template int dispatch_bn<__nv_bfloat16, float>(int, bool, bool, bool, const CUtensorMap&, const CUtensorMap&,
                                                      const CUtensorMap&, const CUtensorMap&, const CUtensorMap&, const GemmArgs&,
                                                      cudaStream_t);

// ------------------------------------------------------------------------------------------------
// host side: tensor maps
// ------------------------------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 2-D row-major tensor [rows][cols] with pitch `ld` elements; box = box_cols x box_rows; 128 B swizzle.
int make_tmap_4d(CUtensorMap* out, const void* ptr, int elem_bytes, bool is_bf16, long long rows, long long cols,
                 long long ld, int box_cols, int box_rows, bool atom32, long long nb0, long long sb0, long long nb1,
                 long long sb1) {
  EncodeTiledFn fn = get_encode_fn();
  if (!fn) return fail(B200_ERR_DRIVER, "cuTensorMapEncodeTiled entry point unavailable");
  if ((reinterpret_cast<uintptr_t>(ptr) & 15) != 0) return fail(B200_ERR_ARG, "TMA base pointer must be 16-byte aligned");
  if (((ld * elem_bytes) & 15) != 0) return fail(B200_ERR_ARG, "TMA row pitch (%lld B) must be a multiple of 16", ld * elem_bytes);
  if (box_cols * elem_bytes > 128 || box_rows > 256) return fail(B200_ERR_ARG, "TMA box too large");
  if (nb0 < 1) nb0 = 1;
  if (nb1 < 1) nb1 = 1;
  if (nb0 == 1 && sb0 <= 0) sb0 = ld * rows;
  if (nb1 == 1 && sb1 <= 0) sb1 = ld * rows;
  if (((sb0 * elem_bytes) & 15) || ((sb1 * elem_bytes) & 15)) return fail(B200_ERR_ARG, "TMA batch strides must be multiples of 16 bytes");
  cuuint64_t dims[4] = {(cuuint64_t)cols, (cuuint64_t)rows, (cuuint64_t)nb0, (cuuint64_t)nb1};
  cuuint64_t strides[3] = {(cuuint64_t)(ld * elem_bytes), (cuuint64_t)(sb0 * elem_bytes), (cuuint64_t)(sb1 * elem_bytes)};
  cuuint32_t box[4] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows, 1, 1};
  cuuint32_t estr[4] = {1, 1, 1, 1};
  CUtensorMapDataType dt = elem_bytes == 4 ? CU_TENSOR_MAP_DATA_TYPE_FLOAT32
                                           : (is_bf16 ? CU_TENSOR_MAP_DATA_TYPE_BFLOAT16 : CU_TENSOR_MAP_DATA_TYPE_FLOAT16);
  CUresult r = fn(out, dt, 4, const_cast<void*>(ptr), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  atom32 ? CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) return fail(B200_ERR_DRIVER, "cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%lld ld=%lld box=%dx%d",
                                     (int)r, rows, cols, ld, box_cols, box_rows);
  return 0;
}


// Pick the N tile: minimise (waves x tile cost); BN=64 re-reads A from smem at twice the tensor-core rate,
// so it carries a penalty and is only chosen for very small problems.
static int choose_bn(int M, int N, int splits, int nbatch) {
  const int sms = num_sms();
  const int cands[4] = {256, 192, 128, 64};
  double best = 1e30;
  int best_bn = 128;
  const int num_m = (M + kBM - 1) / kBM;
  for (int i = 0; i < 4; ++i) {
    const int bn = cands[i];
    const int num_n = (N + bn - 1) / bn;
    if (bn > 64 && bn - (num_n * bn - N) <= bn / 2 && num_n == 1) continue;   // mostly-empty single tile: try narrower
    const long long tiles = (long long)num_m * num_n * splits * nbatch;
    const long long waves = (tiles + sms - 1) / sms;
    double cost = (double)waves * (bn == 64 ? 96.0 : (double)bn) + 8.0 * waves;  // + per-tile fixed overhead
    if (cost < best - 1e-9) {
      best = cost;
      best_bn = bn;
    }
  }
  return best_bn;
}

// Does the CTA-pair kernel pay for this problem?  It halves the B traffic per SM but schedules 256-row tiles on 74
// pairs.  Measured on B200 (profiles/README.md §9): 8-18 % faster on 8192^3 (except fp32 with MN-major A: 17 % slower),
// 4-7 % on [4096 x 3072 x 768], nothing on the narrower BERT-base GEMMs, and a BERT-base training step with it forced
// everywhere is 1.3 % SLOWER (coarser tiles with < 3 waves), a BERT-large step 0.5 % faster.  Hence: >= 3 waves only
// (in BERT-base that is the two [4096 x 3072 x 768] GEMMs: 35.2 -> 33.7 us and 39.1 -> 36.3 us).
static bool cta2_pays(int M, int N, int bn, int splits, int nbatch, bool fp32_a_mn) {
  if (fp32_a_mn) return false;
  const int sms = num_sms();
  const long long t1 = (long long)((M + 127) / 128) * ((N + bn - 1) / bn) * splits * nbatch;
  const long long t2 = (long long)((M + 255) / 256) * ((N + bn - 1) / bn) * splits * nbatch;
  const long long w1 = (t1 + sms - 1) / sms, w2 = (t2 + sms / 2 - 1) / (sms / 2);
  return w2 <= w1 && w2 >= 3;
}

static int gemm_impl(const b200_gemm_desc* d, cudaStream_t stream) {
  if (!d || !d->A || !d->B || !d->D) return fail(B200_ERR_ARG, "gemm: null pointer");
  if (d->M <= 0 || d->N <= 0 || d->K <= 0) return fail(B200_ERR_ARG, "gemm: bad shape %dx%dx%d", d->M, d->N, d->K);
  const bool in_bf16 = d->dtype != B200_F32_TF32;
  const bool out_bf16 = d->dtype == B200_BF16;
  const int es = in_bf16 ? 2 : 4;
  const int os = out_bf16 ? 2 : 4;
  if (d->accumulate && out_bf16) return fail(B200_ERR_UNSUPPORTED, "gemm: accumulate needs fp32 output");
  if (d->aux_mode != B200_AUX_NONE) {
    if (!d->aux) return fail(B200_ERR_ARG, "gemm: aux_mode set but aux is null");
    if ((d->N % (16 / os)) != 0 || ((d->ldaux * os) & 15) != 0 || (reinterpret_cast<uintptr_t>(d->aux) & 15))
      return fail(B200_ERR_ARG, "gemm: aux needs N and pitch multiple of 16 bytes and an aligned base");
  }
  if (d->act == B200_ACT_GELU_SAVE_GRAD && !d->D2) return fail(B200_ERR_ARG, "gemm: GELU_SAVE_GRAD needs D2 for the derivative");
  if (d->bias && (d->N % (16 / os)) == 0 && (reinterpret_cast<uintptr_t>(d->bias) & 15))
    return fail(B200_ERR_ARG, "gemm: bias must be 16-byte aligned");

  const int BK = 128 / es;
  const int EA = 128 / es;
  const int kb_total = (d->K + BK - 1) / BK;
  const int num_m = (d->M + kBM - 1) / kBM;

  int splits = 1;
  int joint_bn = 0;
  if (d->accumulate) {
    splits = d->splits;
    if (splits <= 0) {
      // weight gradients: pick (BN, split-K) together so that the tile count fills whole waves of SMs.
      // cost ~ waves x (k-blocks per split x tile width + epilogue); >= 8 k-blocks per split.
      const int cands[4] = {256, 192, 128, 64};
      double best = 1e30;
      splits = 1;
      const int nb = (d->batch0 > 1 ? d->batch0 : 1) * (d->batch1 > 1 ? d->batch1 : 1);
      for (int ci = 0; ci < 4; ++ci) {
        const int bn = cands[ci];
        if (d->bn > 0 && bn != d->bn) continue;
        const long long base = (long long)num_m * ((d->N + bn - 1) / bn) * nb;
        for (int sp = 1; sp <= 16; ++sp) {
          if (sp > 1 && kb_total / sp < 8) break;
          const long long tiles = base * sp;
          const long long waves = (tiles + num_sms() - 1) / num_sms();
          const int kbs = (kb_total + sp - 1) / sp;
          // cycles per k-block: the slower of the tensor core (~2.87 cycles per output column at the sustained
          // rate) and operand ingest ((128 + bn) rows of 128 B at ~60 B/clk per SM); wide tiles amortise the A
          // tile, which is why 256 x split-2 beats 128 x split-1 by 26 % on the FFN weight gradients
          // (tools/gpu_sweep_wgrad.py, profiles/README.md §10).  Plus the per-tile epilogue (TMEM -> TMA reduce-add).
          const double mma = 2.87 * (bn == 64 ? 96.0 : (double)bn);
          const double ingest = 2.15 * (128.0 + bn);
          const double cost = (double)waves * ((double)kbs * (mma > ingest ? mma : ingest) + 5.0 * bn + 800.0);
          if (cost < best - 1e-9) { best = cost; splits = sp; joint_bn = bn; }
        }
      }
    }
  }
  // Skinny outputs with a long reduction (e.g. the MLM decoder dgrad: 640 x 768 over K = 30522) would leave most
  // SMs idle: zero the output and let several CTAs per tile add their K-slices with the TMA reduce-add store.
  bool auto_split = false;
  int forced_bn = 0;
  if (!d->accumulate && !out_bf16 && d->act == B200_ACT_NONE && d->aux_mode == B200_AUX_NONE && !d->D2 &&
      (d->batch0 <= 1 && d->batch1 <= 1) && d->splits >= 0) {
    const long long tiles128 = (long long)num_m * ((d->N + 127) / 128);
    if (tiles128 * 2 <= num_sms() && kb_total >= 64) {
      int sp = (int)(num_sms() / tiles128);
      if (sp > kb_total / 16) sp = kb_total / 16;
      if (sp > 16) sp = 16;
      if (sp >= 2) {
        splits = sp;
        auto_split = true;
        forced_bn = d->bn > 0 ? d->bn : 128;
        B200_CUDA(cudaMemset2DAsync(d->D, (size_t)d->ldd * os, 0, (size_t)d->N * os, (size_t)d->M, stream));
      }
    }
  }
  int kb_per_split = (kb_total + splits - 1) / splits;
  splits = (kb_total + kb_per_split - 1) / kb_per_split;  // no empty split

  const int nb0 = d->batch0 > 0 ? d->batch0 : 1;
  const int nb1 = d->batch1 > 0 ? d->batch1 : 1;
  const int nbatch = nb0 * nb1;
  const int epi_mode = d->epi_mode;
  if (epi_mode == 0 && nbatch > 1 && (d->bias || d->aux || d->D2))
    return fail(B200_ERR_UNSUPPORTED, "gemm: batched mode has no bias/aux/D2 epilogue");
  if (epi_mode != 0) {
    if (epi_mode != 1 && epi_mode != 2) return fail(B200_ERR_ARG, "gemm: unknown epi_mode %d", epi_mode);
    if (d->N > 256) return fail(B200_ERR_UNSUPPORTED, "gemm: softmax epilogues need the whole row in one tile (N <= 256)");
    if (d->accumulate || d->bias || d->act) return fail(B200_ERR_ARG, "gemm: softmax epilogues take no bias/act/accumulate");
    if (epi_mode == 2 && !d->aux) return fail(B200_ERR_ARG, "gemm: softmax-backward epilogue needs the probabilities as aux");
    if (d->p_drop > 0.f && !d->seed_dev) return fail(B200_ERR_ARG, "gemm: dropout needs seed_dev");
    if (epi_mode == 1 && d->p_drop > 0.f && !d->D2) return fail(B200_ERR_ARG, "gemm: softmax with dropout needs D2 for the dropped copy");
  }
  int bn = forced_bn > 0 ? forced_bn
                         : (d->bn > 0 ? d->bn : (joint_bn > 0 ? joint_bn : choose_bn(d->M, d->N, splits, nbatch)));
  if (epi_mode != 0) bn = d->N <= 64 ? 64 : (d->N <= 128 ? 128 : (d->N <= 192 ? 192 : 256));
  if (bn != 64 && bn != 128 && bn != 192 && bn != 256) return fail(B200_ERR_ARG, "gemm: bn must be 64/128/192/256");
  // CTA pairs (cta_group::2): a CTA ingests A + B/2 per k-block instead of A + B.  B200_GEMM_CTA2 = 0 / 1 forces it.
  bool cta2 = false;
  {
    const char* e = getenv("B200_GEMM_CTA2");      // read per call so that tests can flip it
    const int env = e ? atoi(e) : -1;
    const bool can = epi_mode == 0 && (bn == 128 || bn == 256) && d->M > kBM;
    cta2 = can && (env == 1 || (env < 0 && cta2_pays(d->M, d->N, bn, splits, nbatch, !in_bf16 && d->a_mn)));
  }
  const int tile_m = cta2 ? 2 * kBM : kBM;
  const int num_m_eff = (d->M + tile_m - 1) / tile_m;

  CUtensorMap tA, tB, tD, tD2;
  int rc;
  if (!d->a_mn) rc = make_tmap_4d(&tA, d->A, es, in_bf16, d->M, d->K, d->lda, BK, kBM, false, nb0, d->sa0, nb1, d->sa1);
  else rc = make_tmap_4d(&tA, d->A, es, in_bf16, d->K, d->M, d->lda, EA, BK, es == 4, nb0, d->sa0, nb1, d->sa1);
  if (rc) return rc;
  if (!d->b_mn) rc = make_tmap_4d(&tB, d->B, es, in_bf16, d->N, d->K, d->ldb, BK, cta2 ? bn / 2 : bn, false, nb0, d->sb0, nb1, d->sb1);
  else rc = make_tmap_4d(&tB, d->B, es, in_bf16, d->K, d->N, d->ldb, EA, BK, es == 4, nb0, d->sb0, nb1, d->sb1);
  if (rc) return rc;
  const int CW = 128 / os;
  rc = make_tmap_4d(&tD, d->D, os, out_bf16, d->M, d->N, d->ldd, CW, 32, false, nb0, d->sd0, nb1, d->sd1);
  if (rc) return rc;
  if (d->D2) {
    rc = make_tmap_4d(&tD2, d->D2, os, out_bf16, d->M, d->N, d->ldd, CW, 32, false, nb0, d->sd0, nb1, d->sd1);
    if (rc) return rc;
  } else {
    tD2 = tD;
  }
  CUtensorMap tX = tD;
  if (d->aux) {
    rc = make_tmap_4d(&tX, d->aux, os, out_bf16, d->M, d->N, d->ldaux, CW, 32, false, nb0, d->sd0, nb1, d->sd1);
    if (rc) return rc;
  }

  GemmArgs a;
  a.M = d->M; a.N = d->N; a.K = d->K;
  a.num_m = num_m_eff;
  a.num_n = (d->N + bn - 1) / bn;
  a.splits = splits;
  a.nb0 = nb0;
  a.nbatch = nbatch;
  a.kb_total = kb_total;
  a.kb_per_split = kb_per_split;
  a.bias = d->bias;
  a.aux = d->aux;
  a.ldaux = d->ldaux;
  a.act = d->act;
  a.aux_mode = d->aux_mode;
  a.accumulate = d->accumulate || auto_split;
  a.has_d2 = d->D2 != nullptr;
  a.alpha = d->alpha;
  a.epi_mode = epi_mode;
  a.attn_mask = d->attn_mask;
  a.p_drop = d->p_drop;
  a.seed_dev = d->seed_dev;
  a.stream_id = d->stream_id;
  a.ldp = d->ldd;
  if (epi_mode != 0) a.aux_mode = B200_AUX_NONE;     // aux is consumed by the softmax-backward epilogue itself

  switch (d->dtype) {
    case B200_F32_TF32: return dispatch_bn<float, float>(bn, cta2, d->a_mn, d->b_mn, tA, tB, tD, tD2, tX, a, stream);
    case B200_BF16: return dispatch_bn<__nv_bfloat16, __nv_bfloat16>(bn, cta2, d->a_mn, d->b_mn, tA, tB, tD, tD2, tX, a, stream);
    case B200_BF16_F32OUT: return dispatch_bn<__nv_bfloat16, float>(bn, cta2, d->a_mn, d->b_mn, tA, tB, tD, tD2, tX, a, stream);
  }
  return fail(B200_ERR_ARG, "gemm: unknown dtype %d", d->dtype);
}

}  // namespace b200

extern "C" int b200_gemm(const b200_gemm_desc* d, void* stream) {
  return b200::gemm_impl(d, reinterpret_cast<cudaStream_t>(stream));
}
