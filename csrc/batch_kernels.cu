// batch_kernels.cu — BERT pre-training batch assembly on the device.
//
// The loader ships ONE packed int32 matrix per batch, row r =
//     input_ids[L] | segment_ids[L] | input_mask[L] | masked_lm_positions[P] | masked_lm_ids[P] | next_sentence_label
// exactly as the six shard datasets store it (hetseq/data/h5pyDataset.py:33-40).  This kernel widens it to the int64
// tensors BertForPreTraining.forward takes and builds the dense label row the reference builds per sample in Python
// (h5pyDataset.py:42-48): labels = -1, then labels[positions[:k]] = ids[:k] with k = index of the first zero position.
// H2D traffic drops from 8*(4L+1) to 4*(3L+2P+1) bytes per sample and the host does no per-sample work.
#include "common.h"
#include "../include/b200seq.h"

namespace b200 {

__global__ void __launch_bounds__(128)
assemble_bert_batch_kernel(const int* __restrict__ raw, int L, int P, long long* __restrict__ ids,
                           long long* __restrict__ seg, long long* __restrict__ mask, long long* __restrict__ labels,
                           long long* __restrict__ nsl) {
  const int row = blockIdx.x;
  const int W = 3 * L + 2 * P + 1;
  const int* r = raw + (long long)row * W;
  const long long o = (long long)row * L;
  __shared__ int first_zero;
  if (threadIdx.x == 0) first_zero = P;
  for (int i = threadIdx.x; i < L; i += blockDim.x) {
    ids[o + i] = r[i];
    seg[o + i] = r[L + i];
    mask[o + i] = r[2 * L + i];
    labels[o + i] = -1;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < P; i += blockDim.x)
    if (r[3 * L + i] == 0) atomicMin(&first_zero, i);
  __syncthreads();
  if (threadIdx.x == 0) {
    // serial on purpose: NumPy's fancy assignment lets the LAST duplicate position win, so does this loop
    const int k = first_zero;
    for (int i = 0; i < k; ++i) {
      const int pos = r[3 * L + i];
      if (pos >= 0 && pos < L) labels[o + pos] = r[3 * L + P + i];
    }
    nsl[row] = r[3 * L + 2 * P];
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_assemble_bert_batch(const int* raw, int B, int L, int P, long long* input_ids, long long* segment_ids,
                                        long long* input_mask, long long* masked_lm_labels, long long* next_sentence_labels,
                                        void* stream) {
  if (!raw || !input_ids || !segment_ids || !input_mask || !masked_lm_labels || !next_sentence_labels)
    return fail(B200_ERR_ARG, "assemble_bert_batch: null pointer");
  if (B < 0 || L < 1 || P < 0) return fail(B200_ERR_ARG, "assemble_bert_batch: bad shape");
  if (B == 0) return 0;
  assemble_bert_batch_kernel<<<B, 128, 0, (cudaStream_t)stream>>>(raw, L, P, input_ids, segment_ids, input_mask,
                                                                  masked_lm_labels, next_sentence_labels);
  B200_LAUNCHED("assemble_bert_batch_kernel");
  return 0;
}
