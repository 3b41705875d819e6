"""Sweep (BN, split-K, CTA pair) for the weight-gradient GEMMs of BERT (dW[M_out, N_in] += dY^T X over K = tokens)
and for the other per-layer GEMM shapes; prints the best configurations next to what the library's own choice takes.

    python tools/gpu_sweep_wgrad.py [bf16]
"""
import os
import sys

import torch

sys.path.insert(0, ".")
from hetseq_9cme_b200.ops.gemm import gemm            # noqa: E402


def timeit(fn, iters=10, reps=3):
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        for _ in range(iters):
            fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (iters * reps) * 1e3


def main():
    dt = torch.bfloat16 if len(sys.argv) > 1 and sys.argv[1] == "bf16" else torch.float32
    T = 4096
    flush = torch.empty(64 * 1024 * 1024, device="cuda")         # 256 MB: evict L2 between launches inside the graph
    for (name, Mo, Ni) in [("Wo 768x768", 768, 768), ("Wqkv 2304x768", 2304, 768), ("W1 3072x768", 3072, 768), ("W2 768x3072", 768, 3072)]:
        dy = torch.randn(T, Mo, device="cuda").to(dt)
        x = torch.randn(T, Ni, device="cuda").to(dt)
        acc = torch.zeros(Mo, Ni, device="cuda")
        res = []
        for cta2 in ("0", "1"):
            os.environ["B200_GEMM_CTA2"] = cta2
            for bn in (64, 128, 192, 256):
                if cta2 == "1" and bn not in (128, 256):
                    continue
                for sp in (1, 2, 3, 4, 6, 8):
                    try:
                        t = timeit(lambda: gemm(dy, x, a_mn=True, b_mn=True, out=acc, out_dtype=torch.float32, accumulate=True, bn=bn, splits=sp))
                    except Exception as e:           # unsupported combination
                        continue
                    res.append((t, bn, sp, cta2))
        os.environ.pop("B200_GEMM_CTA2", None)
        auto = timeit(lambda: gemm(dy, x, a_mn=True, b_mn=True, out=acc, out_dtype=torch.float32, accumulate=True))
        res.sort()
        fl = 2.0 * T * Mo * Ni
        print("%-14s %s auto %.1f us (%.0f TF/s) | best: %s" % (
            name, str(dt)[6:], auto, fl / auto / 1e6,
            "  ".join("bn=%d sp=%d cta2=%s %.1f us" % (bn, sp, c, t) for t, bn, sp, c in res[:5])), flush=True)
    # forward (K-major x K-major) and data-gradient (K-major x MN-major) shapes: BN and CTA pairing only
    for (name, N, K, b_mn) in [("fwd QKV", 2304, 768, False), ("fwd proj", 768, 768, False), ("fwd FFN1", 3072, 768, False),
                               ("fwd FFN2", 768, 3072, False), ("dgrad QKV", 768, 2304, True), ("dgrad proj", 768, 768, True),
                               ("dgrad FFN2", 3072, 768, True), ("dgrad FFN1", 768, 3072, True)]:
        a = torch.randn(T, K, device="cuda").to(dt)
        b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(dt)
        out = torch.empty(T, N, device="cuda", dtype=dt)
        res = []
        for cta2 in ("0", "1"):
            os.environ["B200_GEMM_CTA2"] = cta2
            for bn in (64, 128, 192, 256):
                if cta2 == "1" and bn not in (128, 256):
                    continue
                try:
                    res.append((timeit(lambda: gemm(a, b, b_mn=b_mn, out=out, bn=bn)), bn, cta2))
                except Exception:
                    continue
        os.environ.pop("B200_GEMM_CTA2", None)
        auto = timeit(lambda: gemm(a, b, b_mn=b_mn, out=out))
        res.sort()
        print("%-14s %s auto %.1f us (%.0f TF/s) | best: %s" % (name, str(dt)[6:], auto, 2.0 * T * N * K / auto / 1e6,
              "  ".join("bn=%d cta2=%s %.1f us" % (bn, c, t) for t, bn, c in res[:4])), flush=True)


if __name__ == "__main__":
    main()
