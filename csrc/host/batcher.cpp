// batcher.cpp — host-side batch construction (libb200seq_host.so, no CUDA dependency).
//
// Greedy token-budget batching with the cut rules of the reference's Cython helper
// (hetseq/data/data_utils_fast.pyx:21-62), but driven by a sizes ARRAY instead of a Python callback per
// sample (the reference's callback does a bisect per element, which is why its start-up is slow,
// README.md:71).  Batches are always contiguous runs of the ordered index list, so the result is returned
// as cut positions; a constant-size fast path handles the BERT corpora where every sample reports the same
// length.
#include <algorithm>
#include <cstdint>
#include <cstdio>
#include <vector>

extern "C" {

// sizes[i] = token count of the i-th element of the ordered index list.  cuts_out must hold n+1 entries;
// on return cuts_out[0..nb] are batch boundaries (batch b = [cuts[b], cuts[b+1])).  Returns the number of
// batches, or -(i+1) if element i alone exceeds max_tokens.
long long b200_batch_by_size(const int64_t* sizes, long long n, long long max_tokens, long long max_sentences,
                             long long bsz_mult, int64_t* cuts_out) {
  if (n <= 0) {
    cuts_out[0] = 0;
    return 0;
  }
  if (max_tokens <= 0) max_tokens = INT64_MAX;
  if (max_sentences <= 0) max_sentences = INT64_MAX;
  if (bsz_mult <= 0) bsz_mult = 1;

  long long nb = 0;
  cuts_out[0] = 0;
  long long start = 0;       // first element of the batch under construction
  int64_t sample_len = 0;    // max size over [start, i]
  for (long long i = 0; i < n; ++i) {
    const int64_t nt = sizes[i];
    sample_len = std::max(sample_len, nt);
    if (sample_len > max_tokens) return -(i + 1);
    const long long len = i - start;
    const bool full = len > 0 && (len == max_sentences || (len + 1) > max_tokens / std::max<int64_t>(sample_len, 1) ||
                                  (__int128)(len + 1) * sample_len > (__int128)max_tokens);
    if (full) {
      const long long mod_len = std::max(bsz_mult * (len / bsz_mult), len % bsz_mult);
      start += mod_len;
      cuts_out[++nb] = start;
      sample_len = 0;
      for (long long j = start; j <= i; ++j) sample_len = std::max(sample_len, sizes[j]);
    }
  }
  if (start < n) cuts_out[++nb] = n;
  return nb;
}

// out[i] = index of the shard that holds global sample i (cum = cumulative sizes, length nshards), and
// local[i] = offset inside it.  Vectorised replacement for the per-sample bisect of
// hetseq/data/h5pyDataset.py:88-98.
void b200_locate_samples(const int64_t* cum, long long nshards, const int64_t* idx, long long n, int64_t* shard_out,
                         int64_t* local_out) {
  for (long long i = 0; i < n; ++i) {
    const int64_t* it = std::upper_bound(cum, cum + nshards, idx[i]);
    const long long s = it - cum;
    shard_out[i] = s;
    local_out[i] = idx[i] - (s > 0 ? cum[s - 1] : 0);
  }
}

// masked_lm_labels[b][pos] = ids for the leading non-zero positions of each row; -1 elsewhere
// (hetseq/data/h5pyDataset.py:42-48, vectorised over a whole batch).
void b200_scatter_mlm_labels(const int64_t* positions, const int64_t* ids, long long rows, long long P, long long L,
                             int64_t* labels_out) {
  std::fill(labels_out, labels_out + rows * L, (int64_t)-1);
  for (long long r = 0; r < rows; ++r) {
    for (long long j = 0; j < P; ++j) {
      const int64_t pos = positions[r * P + j];
      if (pos == 0) break;   // zero padding terminates the list
      if (pos > 0 && pos < L) labels_out[r * L + pos] = ids[r * P + j];
    }
  }
}

}  // extern "C"
