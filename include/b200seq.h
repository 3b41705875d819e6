/* b200seq.h — C ABI of libb200seq.so (hand-written sm_100a kernels + host runtime).
 *
 * Every entry point takes raw device pointers and a cudaStream_t (passed as void*), launches
 * asynchronously on that stream, allocates nothing unless stated, and returns 0 on success or a
 * negative b200 error / positive cudaError_t code.  b200_last_error() returns a static string for
 * the calling thread's most recent failure.
 *
 * "ref:" comments cite the file:line in the HetSeq reference whose behaviour the entry replaces.
 */
#ifndef B200SEQ_H_
#define B200SEQ_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* b200_last_error(void);
int b200_version(void);
/* number of kernel launches issued through this library by the calling process (for gpu_launches). */
long long b200_launch_count(void);
void b200_reset_launch_count(void);
/* a CUDA-graph replay re-issues launches that were counted once at capture time: add them explicitly */
void b200_add_launch_count(long long n);

/* ------------------------------------------------------------------------------------------------
 * GEMM: D[M,N] = epilogue(alpha * sum_k A(m,k) * B(n,k))      tcgen05.mma + TMEM + TMA, persistent.
 * ref: every nn.Linear / F.linear in hetseq/bert_modeling.py:166-172,352-354,388,424,515,526,544
 * ---------------------------------------------------------------------------------------------- */
enum { B200_F32_TF32 = 0, /* fp32 operands fed to kind::tf32, fp32 out */
       B200_BF16 = 1,     /* bf16 operands, bf16 out */
       B200_BF16_F32OUT = 2 /* bf16 operands, fp32 out (weight gradients) */ };
enum { B200_ACT_NONE = 0, B200_ACT_GELU = 1, B200_ACT_TANH = 2,
       B200_ACT_GELU_SAVE_GRAD = 3 /* D = gelu(x), D2 = gelu'(x): backward then only multiplies (B200_AUX_MUL) */ };
enum { B200_AUX_NONE = 0, B200_AUX_ADD = 1, /* D = f(acc) + aux                          */
       B200_AUX_MUL_DGELU = 2,              /* D = acc * gelu'(aux), aux = pre-activation */
       B200_AUX_MUL_DTANH = 3, /* D = acc * (1 - aux^2), aux = tanh output */
       B200_AUX_MUL = 4 /* D = acc * aux (aux = a stored activation derivative) */ };

typedef struct {
  int dtype;
  int M, N, K;
  const void* A; long long lda; int a_mn; /* a_mn=0: A is [M][K] (K contiguous); 1: A is [K][M] (M contiguous) */
  const void* B; long long ldb; int b_mn; /* b_mn=0: B is [N][K] (K contiguous); 1: B is [K][N] (N contiguous) */
  void* D; long long ldd;                 /* [M][N], N contiguous */
  void* D2;                               /* optional second output: value before the activation (ldd pitch) */
  const void* bias;                       /* optional [N], output dtype */
  const void* aux; long long ldaux;       /* optional [M][N], output dtype */
  int act, aux_mode;
  int accumulate;                         /* 1: D += result (TMA reduce-add at L2; fp32 out only) */
  float alpha;
  int bn;                                 /* 0 = auto; else 64/128/192/256 */
  int splits;                             /* 0 = auto split-K (only when accumulate=1) */
  /* batched mode (attention): batch index = b1 * batch0 + b0; strides in elements; 0/1 = not batched.
   * No bias / aux / D2 epilogue in batched mode. */
  int batch0, batch1;
  long long sa0, sa1, sb0, sb1, sd0, sd1;
  /* attention epilogues (N <= 256 so that a row lives in one tile; D2 / aux share D's pitch and batch strides):
   *   epi_mode 1: D = softmax(alpha*acc + (1-mask)*-10000) per row, D2 = dropout(D) when p_drop > 0
   *   epi_mode 2: D = P * (dPe - rowsum(P * dPe)),  P = aux,  dPe = acc through the regenerated dropout mask
   * ref: hetseq/bert_modeling.py:361-371 and its autograd dual */
  int epi_mode;
  const long long* attn_mask;            /* [batch1][N] 0/1, optional */
  float p_drop;
  const unsigned long long* seed_dev;
  unsigned long long stream_id;
} b200_gemm_desc;

int b200_gemm(const b200_gemm_desc* d, void* stream);


/* ------------------------------------------------------------------------------------------------
 * Flat-arena optimizer step (all parameters / gradients / moments live in single fp32 buffers).
 * ref: hetseq/optim.py:59-70 (multiply_grads, clip_grad_norm), :162-231 (Adam.step), :263-304
 *      (Adadelta.step); hetseq/controller.py:338-363 (scale -> clip -> step with host syncs)
 * ---------------------------------------------------------------------------------------------- */
/* out[0] (=|+=) sum_i g[i]^2, deterministic.  `scratch` is caller-owned device memory of at least
 * b200_sumsq_scratch_bytes() bytes whose last 4 bytes (arrival counter) must be zero before the first
 * use; the kernel leaves the counter at zero again.  Launches sharing a scratch must not overlap. */
long long b200_sumsq_scratch_bytes(void);
int b200_sumsq_f32(const float* g, long long n, float* out, void* scratch, long long scratch_bytes,
                   int accumulate, void* stream);

typedef struct {
  float lr, beta1, beta2, eps, weight_decay;
  int step;                  /* 1-based update count (bias correction) */
  float grad_scale_mult;     /* gradient scale = mult / *sample_size (or mult when sample_size is NULL) */
  float max_norm;            /* <= 0: no clipping */
  const float* norm_sq;      /* device: sum g^2 of the unscaled gradient; NULL: no clip, gnorm = 0 */
  const double* sample_size; /* device: optional denominator */
  float* stats_out;          /* device (or mapped pinned) float[3]: gnorm, applied scale, clipped flag */
  int zero_grad;             /* 1: leave g zeroed for the next accumulation */
  const float* sample_size_f32; /* device: fp32 alternative to sample_size */
  const float* hyper_dev;    /* device float[4] {lr, lr*sqrt(1-b2^t)/(1-b1^t), 1 - wd*lr, -}: when set, these override
                              * lr/step so that a captured CUDA graph can be replayed with a new schedule value */
} b200_adam_args;
int b200_adam_flat(float* p, float* g, float* m, float* v, void* p_bf16 /* optional bf16 mirror */, long long n,
                   const b200_adam_args* a, void* stream);

typedef struct {
  float lr, rho, eps, weight_decay;
  float grad_scale_mult, max_norm;
  const float* norm_sq;
  const double* sample_size;
  float* stats_out;
  int zero_grad;
  const float* sample_size_f32;
  const float* hyper_dev;    /* device float[4] {lr, -, -, -} */
} b200_adadelta_args;
int b200_adadelta_flat(float* p, float* g, float* square_avg, float* acc_delta, void* p_bf16, long long n,
                       const b200_adadelta_args* a, void* stream);

/* x *= c * (c_dev ? *c_dev : 1) */
int b200_scale_f32(float* x, long long n, float c, const float* c_dev, void* stream);
int b200_cast_f32_to_bf16(const float* src, void* dst_bf16, long long n, void* stream);
/* Operand split of the strict-fp32 GEMM policy (3xTF32): hi = x with the mantissa bits a TF32 operand ignores
 * cleared, lo = the remainder cut the same way; A.B ~= Ahi.Bhi + Ahi.Blo + Alo.Bhi to about 2^-22 (the reference
 * computes in fp32: hetseq/options.py:11).  16-byte aligned buffers. */
int b200_split_tf32(const float* x, float* hi, float* lo, long long n, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Fused LayerNorm:  y = drop_post( LN( drop_pre(x + bias) + residual ) * gamma + beta )
 * ref: hetseq/bert_modeling.py:276-289 (BertLayerNorm, eps inside sqrt), :387-391, :423-427 (dropout +
 *      residual + LN after the attention-output / FFN-2 projections), :318-319 (embedding LN + dropout)
 * dtype: 0 = fp32 tensors, 1 = bf16 tensors (statistics and parameter gradients are always fp32).
 * Dropout masks are regenerated from *seed_dev + stream id + element index; nothing is stored.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  int dtype, rows, cols;
  const void* x; long long ldx;
  const void* bias;                         /* optional [cols] */
  const void* residual; long long ldr;      /* optional [rows][cols] */
  const void* gamma; const void* beta;      /* [cols] */
  float eps, p_pre, p_post;
  const unsigned long long* seed_dev;       /* device uint64, required when a dropout prob > 0 */
  unsigned long long stream_pre, stream_post;
  void* y; long long ldy;                   /* fwd out */
  float* mean; float* rstd;                 /* fwd out / bwd in, [rows] */
  /* backward only */
  const void* dy; long long lddy;
  void* ds; long long ldds;                 /* optional: grad wrt the LN input (== grad of residual) */
  void* dz; long long lddz;                 /* optional: grad wrt x (ds through the pre-dropout mask) */
  float* dgamma; float* dbeta; float* dbias;/* fp32 [cols], accumulated with +=; dbias optional */
  void* scratch; long long scratch_bytes;   /* >= b200_ln_bwd_scratch_bytes(cols); contents need no init.  May be NULL
                                             * when cols % 256 == 0: block partials then go into dgamma/dbeta/dbias with
                                             * fp32 atomics (one launch less, summation order varies) */
} b200_ln_args;
long long b200_ln_bwd_scratch_bytes(int cols);
int b200_ln_fwd(const b200_ln_args* a, void* stream);
int b200_ln_bwd(const b200_ln_args* a, void* stream);
/* Streaming (cp.async) LayerNorm kernels: pipe bit 0 = forward, bit 1 = backward; *_ctas = CTAs per SM of their
 * grids (1 or 2).  pipe = -1: automatic (streaming backward for fp32 storage); pipe < -1 / ctas <= 0 leave the field
 * unchanged.  Returns the mode in effect.  Start-up values: B200_LN_PIPE, B200_LN_PIPE_FWD_CTAS, B200_LN_PIPE_BWD_CTAS. */
int b200_ln_config(int pipe, int fwd_ctas, int bwd_ctas);

/* ------------------------------------------------------------------------------------------------
 * Embeddings, masked-row compaction, cross-entropy, tiny-N heads, column sums.
 * dtype: 0 = fp32 tensors, 1 = bf16 tensors.  Parameter gradients are always fp32 and accumulated (+=).
 * ref: hetseq/bert_modeling.py:306-320 (BertEmbeddings), :899-905 (MLM + NSP CrossEntropyLoss),
 *      :576-581 (heads), :1221-1245 (token-classification head + masked CE)
 * ---------------------------------------------------------------------------------------------- */
/* out[r] = W[ids[r]] + P[r % L] + TT[tts[r]] */
int b200_embed_sum(int dtype, const long long* ids, const long long* tts, const void* W, const void* P, const void* TT,
                   void* out, int rows, int L, int H, int vocab, int ntypes, void* stream);
/* gW[ids[r]] += ds[r]; gP[r % L] += ds[r]; gT[tts[r]] += ds[r]   (fp32 vector reductions at L2) */
int b200_embed_scatter(int dtype, const long long* ids, const long long* tts, const void* ds, float* gW, float* gP,
                       float* gT, int rows, int L, int H, int vocab, int ntypes, void* stream);
/* order-preserving compaction of rows with labels != ignore (and mask != 0): idx_out/label_out have `cap`
 * entries (tail = -1 / ignore); count_out = {min(n,cap), n} */
int b200_compact_rows(const long long* labels, const long long* mask, int rows, int ignore, int cap, int* idx_out,
                      long long* label_out, int* count_out, void* stream);
int b200_gather_rows(int dtype, const void* x, long long ldx, const int* idx, int n, int H, void* out, void* stream);
int b200_scatter_rows(int dtype, const void* dy, const int* idx, int n, int H, void* dx, long long lddx, int accumulate,
                      void* stream);
/* wide softmax cross-entropy: per-row logsumexp + loss; backward overwrites logits with d(logits) */
int b200_ce_fwd(int dtype, const void* logits, long long ld, const long long* labels, int rows, int V, int ignore,
                float* lse, float* row_loss, void* stream);
int b200_ce_bwd(int dtype, void* logits, long long ld, const long long* labels, int rows, int V, int ignore,
                const float* lse, const float* gout, const float* inv_count, void* stream);
/* loss = sum(row_loss)/max(#valid,1); inv_count = 1/max(#valid,1) */
int b200_ce_finalize(const float* row_loss, const long long* labels, int rows, int ignore, float* loss, float* inv_count,
                     float* count_out, void* stream);
/* tiny heads (1 <= N <= 32): fp32 logits */
int b200_head_fwd(int dtype, const void* x, long long ldx, const void* w, const void* b, int rows, int K, int N, float* out,
                  void* stream);
int b200_head_ce(const float* logits, const long long* labels, const long long* active, int rows, int N, int ignore,
                 float* row_loss, long long* eff_labels, void* stream);
int b200_head_ce_bwd(const float* logits, const long long* eff_labels, int rows, int N, int ignore, const float* gout,
                     const float* inv_count, float* dlogits, void* stream);
int b200_head_bwd(int dtype, const float* dl, const void* x, long long ldx, const void* w, int rows, int K, int N, void* dx,
                  long long lddx, int accumulate_dx, float* dw, float* db, void* stream);
/* out = dy * act'(aux); mode 1: gelu' at pre-activation aux, mode 2: tanh' = 1 - aux^2 at output aux */
int b200_act_bwd(int dtype, const void* dy, const void* aux, void* out, long long n, int mode, void* stream);
/* y = x * keep_mask/(1-p) (n % 4 == 0); the same call on a gradient is the backward */
int b200_dropout(int dtype, const void* x, void* y, long long n, float p, const unsigned long long* seed_dev,
                 unsigned long long stream_id, void* stream);
/* out[c] += sum_r x[r][c]; scratch >= b200_colsum_scratch_bytes(rows, cols), no init needed (deterministic
 * two-pass).  scratch == NULL selects a single-pass variant that adds block partials to out[] with fp32
 * reductions at L2 (order of additions varies from run to run). */
long long b200_colsum_scratch_bytes(int rows, int cols);
int b200_colsum(int dtype, const void* x, long long ldx, int rows, int cols, float* out, void* scratch,
                long long scratch_bytes, void* stream);
/* Same result as b200_colsum (out[c] += sum_r x[r][c]) for callers that cannot allocate device memory — for
 * instance a host application whose own pool already owns the whole device, or a loop that cannot take the
 * device-wide stall of cudaMalloc/cudaFree — and therefore have no scratch to pass.  This entry point has no
 * counterpart in the reference.  The partial sums go to a 1 MiB array built into the library image (resident on
 * every GPU that loads the library, until exit); cols must be a multiple of 4 and at most
 * b200_colsum_static_max_cols() (= 4096), larger inputs are refused with an error.  Allocates nothing.  The first
 * call on each device resolves the array's device address (make that call before any stream capture begins).
 * All calls share the one array and it keeps the previous call's partials (none of them needs clearing: every
 * launch rewrites what it reads), so calls that may overlap on the GPU race — keep them on one stream or order
 * them with events.  `stream` must belong to the device that is current at the call (the array's address is
 * looked up for the current device).  As for b200_colsum, x must be 16-byte aligned for fp32 (8-byte for bf16)
 * and ldx a multiple of 4 elements and >= cols; violations are refused with an error code. */
int b200_colsum_static_max_cols(void);
int b200_colsum_static(int dtype, const void* x, long long ldx, int rows, int cols, float* out, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Attention softmax over scores S[B][heads][L][Lp] produced by the batched GEMM (already scaled):
 * P = softmax(S + (1-mask)*-10000) written over S; Pdrop = dropout(P) (only when p_drop > 0).
 * Backward: dP <- P * (dPe - rowsum(P*dPe)), dPe = dP through the regenerated dropout mask.
 * ref: hetseq/bert_modeling.py:361-371 (scores/sqrt(d) + mask, Softmax, Dropout), :817-825 (mask)
 * ---------------------------------------------------------------------------------------------- */
int b200_softmax_fwd(int dtype, void* S, void* Pdrop, const long long* attn_mask /* [B][L] 0/1, optional */, int B,
                     int heads, int L, int Lp, float p_drop, const unsigned long long* seed_dev,
                     unsigned long long stream_id, void* stream);
int b200_softmax_bwd(int dtype, const void* P, void* dP, int B, int heads, int L, int Lp, float p_drop,
                     const unsigned long long* seed_dev, unsigned long long stream_id, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Batch assembly on the device (ref: hetseq/data/h5pyDataset.py:33-48 does this per sample on the host).
 * raw[B][3L+2P+1] int32 = input_ids | segment_ids | input_mask | masked_lm_positions | masked_lm_ids | nsp label;
 * outputs are the int64 [B][L] / [B] tensors of BertForPreTraining.forward, labels = -1 except at the positions
 * before the first zero position.
 * ---------------------------------------------------------------------------------------------- */
int b200_assemble_bert_batch(const int* raw, int B, int L, int P, long long* input_ids, long long* segment_ids,
                             long long* input_mask, long long* masked_lm_labels, long long* next_sentence_labels,
                             void* stream);

/* ------------------------------------------------------------------------------------------------
 * Fused multi-head attention, head size 64 (tcgen05 + TMEM + TMA; no [B,heads,L,L] tensor in HBM).
 *   forward : out[B*L, H] = dropout(softmax(Q K^T * scale + (1-mask)*-10000)) V, lse[B][heads][L] saved
 *   backward: out = dqkv[B*L, 3H] from qkv, o (forward output), dout, lse; delta[B][heads][L] is scratch
 * qkv is the packed projection [B*L, 3H] (Q | K | V sections, head-major inside each), H = heads*64.
 * Dropout elements are indexed ((b*heads+h)*L + q)*Lp + k like b200_softmax_fwd, so for one seed/stream id the
 * fused and unfused paths drop the same probabilities.
 * ref: hetseq/bert_modeling.py:341-376 (BertSelfAttention.forward) and its autograd dual
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  int dtype;                         /* 0 = fp32 storage / tf32 math, 1 = bf16 */
  int B, heads, L, Lp;               /* Lp: row pitch used for dropout indexing (>= L, multiple of 4) */
  float scale, p_drop;
  const void* qkv; long long ld_qkv;
  const long long* attn_mask;        /* [B][L] 0/1, optional */
  const unsigned long long* seed_dev;
  unsigned long long stream_id;
  void* out; long long ld_out;
  float* lse;
  float* delta;                      /* backward only */
  const void* o;                     /* backward only */
  const void* dout;                  /* backward only */
  long long ld_o;                    /* pitch of o and dout */
} b200_attn_args;
int b200_attn_fwd(const b200_attn_args* args, void* stream);
int b200_attn_bwd(const b200_attn_args* args, void* stream);

/* ------------------------------------------------------------------------------------------------
 * Peer-to-peer gradient all-reduce over NVLink (replaces DDP's Reducer + ncclAllReduce,
 * ref: hetseq/controller.py:79-87) and the CUDA-IPC plumbing that maps the peers' arenas.
 * ---------------------------------------------------------------------------------------------- */
int b200_ipc_export(const void* ptr, unsigned char* handle_out /* 64 bytes */, long long* offset_out);
int b200_ipc_open(const unsigned char* handle /* 64 bytes */, long long offset, void** ptr_out, void** base_out);
int b200_ipc_close(void* base);
/* data_ptrs[r] / flag_ptrs[r]: rank r's gradient arena / flag array as mapped in THIS process (own rank = local
 * pointer).  Flag arrays hold 2 * max_buckets * 8 uint32, zero-initialised; `counter` holds max_buckets uint32,
 * zero-initialised, local.  `epoch_dev` holds max_buckets uint32 (zero-initialised, local): the kernel derives
 * the flag epoch from it and advances it, so all ranks must issue the same sequence of reductions.  SUM over ranks of arena[start, start+n) ends up in every rank's arena; one_shot=1 (small buckets) reads
 * whole peer buckets, one_shot=0 is reduce-scatter + all-gather.  The kernel spins on peer flags: all ranks must
 * launch it, each on its own GPU. */
int b200_p2p_allreduce(void* const* data_ptrs, void* const* flag_ptrs, void* counter, int rank, int world,
                       long long start, long long n, int bucket, int max_buckets, void* epoch_dev, int one_shot,
                       int blocks, float* sq_out /* optional local fp32: += sum of squares of what this rank reduced */,
                       void* multicast_base /* optional NVLS multicast address of the arena: multimem.ld_reduce / .st */,
                       void* stream);

/* Reduce-scatter half of the above (sharded-optimizer mode): the SUM over ranks of this rank's 1/world slice of
 * arena[start, start+n) is written into the LOCAL arena only (plus sum of squares into sq_out); nothing is sent
 * back.  One flag round per bucket: a later collective on the same ranks (b200_p2p_allreduce of the statistics
 * tail) is what tells a rank that its peers are done reading its gradients.  Slice r of a bucket covers float4
 * indices [r*per, min((r+1)*per, n/4)) with per = ceil(n/4 / world) — b200_p2p_sharded_adam uses the same rule. */
int b200_p2p_reduce_scatter(void* const* data_ptrs, void* const* flag_ptrs, void* counter, int rank, int world,
                            long long start, long long n, int bucket, int max_buckets, void* epoch_dev, int blocks,
                            float* sq_out, void* multicast_base, void* stream);

/* Optimizer step of the sharded mode — the fused reduce -> scale/clip -> Adam -> parameter broadcast path
 * (ref: hetseq/controller.py:338-347 multiply_grads / clip_grad_norm / optimizer.step, hetseq/optim.py:206-229).
 * Each rank applies the reference Adam update to ITS slice of every bucket (ranges[2*b], ranges[2*b+1] = start and
 * length of bucket b in elements, device array, n_buckets entries) using its local, already reduced gradient, and
 * writes the new parameters into every rank's copy: fp32 masters through param_ptrs / param_mc when bf16_ptrs is
 * null, otherwise the fp32 master stays local (only the owner's slice is kept current) and the bf16 working copy
 * is what travels (bf16_ptrs / bf16_mc).  It also zeroes the whole local gradient arena [0, total).  A flag round
 * (slot `bucket`) ends the kernel: when it retires every peer's writes into this rank's parameters have landed.
 * The update may be issued bucket by bucket (ranges + 2*b, n_buckets = 1, a flag slot per bucket) on a side stream
 * so that the next forward pass starts as soon as the first layers' parameters are in.
 * Scalars as in b200_adam_flat (a->norm_sq / sample_size / hyper_dev / stats_out). */
int b200_p2p_sharded_adam(void* const* param_ptrs, void* param_mc, void* const* bf16_ptrs, void* bf16_mc,
                          float* grad, float* m, float* v, const long long* ranges, int n_buckets, long long total,
                          void* const* flag_ptrs, void* counter, int rank, int world, int bucket, int max_buckets,
                          void* epoch_dev, const b200_adam_args* a, int blocks,
                          const float* gate /* optional device float: 0 skips the update, the flag round still runs */,
                          void* stream);

/* How long a kernel waits for a peer's flag before it gives up (seconds; default 600), and an optional word of
 * mapped pinned host memory that then receives (0x80000000 | slot << 8 | missing rank).  A tripped watchdog does
 * not trap: the kernel finishes on incomplete data, later waits return immediately, and the host is expected to
 * check the word after each step and stop.  b200_p2p_watchdog_reset clears the tripped state (tests). */
int b200_p2p_set_watchdog(double timeout_seconds, unsigned int* host_error_word_devptr);
int b200_p2p_watchdog_reset(void);

/* ------------------------------------------------------------------------------------------------
 * Communication runtime (csrc/comm_runtime.cu): rendezvous over Unix sockets, symmetric buffers from the CUDA
 * virtual-memory API mapped into every rank, NVLS multicast mappings.  Replaces what the reference gets from
 * torch.distributed's process group + DDP's bucket allocation (hetseq/distributed_utils.py:11-41,
 * hetseq/controller.py:79-87).  One NVSwitch domain, <= 8 ranks, one process per GPU, device current at init.
 * ---------------------------------------------------------------------------------------------- */
typedef struct {
  void* handle;        /* pass to b200_comm_free */
  long long bytes;     /* mapped size (>= requested, rounded to the allocation granularity) */
  void* local;         /* this rank's buffer, zero-filled */
  void* peer[8];       /* rank r's buffer as mapped in this process (peer[rank] == local) */
  void* multicast;     /* NVLS address covering all ranks' buffers, or NULL */
} b200_comm_buffer;
int b200_comm_init(int rank, int world, const char* tag, int timeout_ms, void** comm_out);
int b200_comm_multicast_supported(void* comm);
int b200_comm_barrier(void* comm);
/* Control-plane self-test without a GPU: rendezvous, agreement, descriptor passing, out-of-order rounds. */
int b200_comm_selftest(int rank, int world, const char* tag, int timeout_ms);
int b200_comm_alloc(void* comm, long long bytes, int want_multicast, b200_comm_buffer* out);   /* collective */
int b200_comm_free(void* comm, void* handle);
int b200_comm_destroy(void* comm);

#ifdef __cplusplus
}
#endif
#endif /* B200SEQ_H_ */
