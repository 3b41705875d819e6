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

/* ------------------------------------------------------------------------------------------------
 * GEMM: D[M,N] = epilogue(alpha * sum_k A(m,k) * B(n,k))      tcgen05.mma + TMEM + TMA, persistent.
 * ref: every nn.Linear / F.linear in hetseq/bert_modeling.py:166-172,352-354,388,424,515,526,544
 * ---------------------------------------------------------------------------------------------- */
enum { B200_F32_TF32 = 0, /* fp32 operands fed to kind::tf32, fp32 out */
       B200_BF16 = 1,     /* bf16 operands, bf16 out */
       B200_BF16_F32OUT = 2 /* bf16 operands, fp32 out (weight gradients) */ };
enum { B200_ACT_NONE = 0, B200_ACT_GELU = 1, B200_ACT_TANH = 2 };
enum { B200_AUX_NONE = 0, B200_AUX_ADD = 1, /* D = f(acc) + aux                          */
       B200_AUX_MUL_DGELU = 2,              /* D = acc * gelu'(aux), aux = pre-activation */
       B200_AUX_MUL_DTANH = 3 /* D = acc * (1 - aux^2), aux = tanh output */ };

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
} b200_gemm_desc;

int b200_gemm(const b200_gemm_desc* d, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* B200SEQ_H_ */
