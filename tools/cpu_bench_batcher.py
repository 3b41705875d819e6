#!/usr/bin/env python3
"""Host-side batch construction: the reference's Cython `batch_by_size_fast` (one Python callback per sample,
hetseq/data/data_utils_fast.pyx) against this repository's C++ batcher fed with a sizes array
(csrc/host/batcher.cpp through data/data_utils.py).  CPU only.  usage: python tools/cpu_bench_batcher.py [n_samples]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from hetseq_9cme_b200.data import data_utils  # noqa: E402


def best(fn, reps=3):
    t = []
    for _ in range(reps):
        t0 = time.perf_counter()
        out = fn()
        t.append(time.perf_counter() - t0)
    return min(t), out


def main():
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_000_000
    rng = np.random.RandomState(0)
    idx = np.arange(n, dtype=np.int64)
    cases = [("BERT shards: constant size 128, 32 sentences / batch", np.full(n, 128, dtype=np.int64), None, 32, 1),
             ("variable length 4-60 tokens, 4096 tokens / batch, multiple of 8", rng.randint(4, 60, size=n).astype(np.int64), 4096, 10 ** 9, 8)]
    try:
        sys.path[:0] = [os.path.join(ROOT, "baseline", "_ref"), os.path.join(ROOT, "baseline", "shims")]   # h5py / boto3 stand-ins
        from hetseq.data.data_utils_fast import batch_by_size_fast as ref_fast
    except Exception as e:  # noqa: BLE001
        ref_fast = None
        print("reference Cython batcher not importable (%s): timing ours only" % type(e).__name__)
    from hetseq_9cme_b200.ops import native
    print("%d samples, best of 3, host lib %s" % (n, "loaded" if native.host_lib() is not None else "MISSING (numpy fallback)"))
    for name, sizes, mt, ms, mult in cases:
        t_ours, got = best(lambda: data_utils.batch_by_size(idx, sizes=sizes, max_tokens=mt, max_sentences=ms, required_batch_size_multiple=mult))
        line = "%-66s ours %.3f s (%d batches)" % (name, t_ours, len(got))
        if ref_fast is not None:
            t_ref, want = best(lambda: ref_fast(idx, lambda i: int(sizes[i]), mt if mt is not None else sys.maxsize, ms, mult), reps=1)
            same = len(want) == len(got) and all(list(map(int, a)) == b for a, b in zip(want[:1000], got[:1000]))
            line += "   reference %.3f s   x%.1f   first 1000 batches %s" % (t_ref, t_ref / t_ours, "identical" if same else "DIFFER")
        print(line, flush=True)


if __name__ == "__main__":
    main()
