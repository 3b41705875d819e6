"""Device time of the fused LayerNorm kernels, direct-load vs cp.async-pipelined (b200_ln_config), at the shapes of
the benchmark step.  24 calls over 12 rotating buffer sets (> L2) in one CUDA graph, so neither host launch cost nor
L2 residency of the inputs flatters the number.  Usage: python tools/gpu_probe_ln.py [--rows 4096 --cols 768]"""
import argparse
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hetseq_9cme_b200.ops import kernels as K  # noqa: E402


def graph_time(fn, calls, reps=10):
    s = torch.cuda.Stream()
    with torch.cuda.stream(s):
        for i in range(calls):
            fn(i)
        torch.cuda.synchronize()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for i in range(calls):
                fn(i)
        for _ in range(3):
            g.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            g.replay()
        e1.record()
        torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1000.0 / (reps * calls)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rows", type=int, default=4096)
    ap.add_argument("--cols", type=int, default=768)
    ap.add_argument("--dtype", default="fp32")
    a = ap.parse_args()
    dt = torch.float32 if a.dtype == "fp32" else torch.bfloat16
    ws = K.Workspace.get("cuda")
    ws.set_seed(5)
    nset, calls = 12, 24
    X = [torch.randn(a.rows, a.cols, device="cuda").to(dt) for _ in range(nset)]
    R = [torch.randn(a.rows, a.cols, device="cuda").to(dt) for _ in range(nset)]
    DY = [torch.randn(a.rows, a.cols, device="cuda").to(dt) for _ in range(nset)]
    bias = torch.randn(a.cols, device="cuda").to(dt)
    gamma, beta = torch.ones(a.cols, device="cuda").to(dt), torch.zeros(a.cols, device="cuda").to(dt)
    dg, db, dbias = (torch.zeros(a.cols, device="cuda") for _ in range(3))
    _, mean, rstd = K.ln_fwd(X[0], gamma, beta, bias=bias, residual=R[0], p_pre=0.1, seed=ws.seed, stream_pre=3)

    def fwd(i):
        K.ln_fwd(X[i % nset], gamma, beta, bias=bias, residual=R[i % nset], p_pre=0.1, seed=ws.seed, stream_pre=3)

    def bwd(i):
        K.ln_bwd(DY[i % nset], X[i % nset], gamma, beta, mean, rstd, dg, db, bias=bias, dbias=dbias, residual=R[i % nset],
                 p_pre=0.1, seed=ws.seed, stream_pre=3, want_dz=True)

    before = K.ln_config()
    nbytes_f = a.rows * a.cols * X[0].element_size() * 3
    nbytes_b = a.rows * a.cols * X[0].element_size() * 5
    print("LayerNorm %d x %d %s, dropout 0.1 + bias + residual; us per call (GB/s)" % (a.rows, a.cols, a.dtype))
    for mode, fc, bc, name in ((0, 2, 2, "direct loads"), (3, 2, 2, "cp.async, 2 CTAs/SM"), (3, 1, 1, "cp.async, 1 CTA/SM"),
                               (3, 2, 1, "cp.async, fwd 2 / bwd 1"), (3, 1, 2, "cp.async, fwd 1 / bwd 2")):
        K.ln_config(mode, fc, bc)
        tf, tb = graph_time(fwd, calls), graph_time(bwd, calls)
        print("  %-26s fwd %6.2f (%5.0f)   bwd %6.2f (%5.0f)" % (name, tf, nbytes_f / tf / 1e3, tb, nbytes_b / tb / 1e3), flush=True)
    K.ln_config(before, 2, 2)
    print("LN PROBE DONE")


if __name__ == "__main__":
    main()
