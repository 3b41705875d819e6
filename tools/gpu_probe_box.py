#!/usr/bin/env python3
"""Times the forward-layout (K-major x K-major) BERT GEMMs with the operand tiles fetched as one TMA box per tile or
as several 64- / 128-row boxes (B200_GEMM_BOX), L2 flushed between launches; plus a correctness check per setting."""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from hetseq_9cme_b200.ops import gemm as G


def timed(fn, flush, n=7):
    for _ in range(3):
        fn()
    ts = []
    for _ in range(n):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    return min(ts) * 1e3


def main():
    flush = torch.empty(256 * 1024 * 1024 // 4, device="cuda")
    shapes = [(4096, 2304, 768), (4096, 768, 768), (4096, 3072, 768), (4096, 768, 3072), (4096, 4096, 1024), (4096, 1024, 4096)]
    for dtype in (torch.bfloat16, torch.float32):
        for (M, N, K) in shapes:
            a = torch.randn(M, K, device="cuda").to(dtype)
            b = torch.randn(N, K, device="cuda").to(dtype)
            bt = b.t().contiguous()
            out = torch.empty(M, N, device="cuda", dtype=dtype)
            ref = (a.double() @ b.double().t())
            res = []
            for box in ("0", "128", "64"):
                os.environ["B200_GEMM_BOX"] = box
                G.gemm(a, b, out=out)
                torch.cuda.synchronize()
                err = (out.double() - ref).abs().max().item() / ref.abs().max().item()
                res.append("box %3s: %6.1f us (err %.1e)" % (box if box != "0" else "one", timed(lambda: G.gemm(a, b, out=out), flush), err))
            os.environ["B200_GEMM_BOX"] = "0"
            t_kmn = timed(lambda: G.gemm(a, bt, b_mn=True, out=out), flush)
            print("%s M=%d N=%d K=%d kk | %s | kmn %.1f us" % ("bf16" if dtype == torch.bfloat16 else "tf32", M, N, K, " | ".join(res), t_kmn), flush=True)


if __name__ == "__main__":
    main()
