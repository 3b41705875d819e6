"""Time the fused tcgen05 attention (csrc/attention_sm100.cu) against the unfused path
(score GEMM -> softmax kernel -> context GEMM, and its 4-GEMM backward) on the shapes the benchmarks use.

    python tools/gpu_probe_attn.py            # prints one line per (shape, dtype)
"""
import math
import sys

import torch

sys.path.insert(0, ".")
from hetseq_9cme_b200.ops import kernels as K          # noqa: E402
from hetseq_9cme_b200.ops.gemm import bgemm            # noqa: E402


def timeit(fn, iters=20, warm=3, reps=5):
    """Device time per call: `iters` calls captured into one CUDA graph (no host launch cost in the number)."""
    for _ in range(warm):
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
    return e0.elapsed_time(e1) / (iters * reps) * 1e3          # microseconds


def main():
    ws = K.Workspace.get("cuda")
    ws.set_seed(1)
    p = 0.1
    for (B, heads, L) in [(32, 12, 128), (8, 12, 512), (32, 16, 128), (4, 16, 512)]:
        for dt in (torch.float32, torch.bfloat16):
            dh, H = 64, heads * 64
            Lp = (L + 7) // 8 * 8
            scale = 1.0 / math.sqrt(dh)
            qkv = torch.randn(B * L, 3 * H, device="cuda").to(dt)
            dctx = torch.randn(B * L, H, device="cuda").to(dt)
            mask = torch.ones(B, L, dtype=torch.int64, device="cuda")
            S = torch.empty(B, heads, L, Lp, device="cuda", dtype=dt)
            dP = torch.empty_like(S)
            ctxv = torch.empty(B * L, H, device="cuda", dtype=dt)
            dqkv = torch.empty_like(qkv)
            sS, sQ, sC = (L * Lp, heads * L * Lp), (dh, L * 3 * H), (dh, L * H)
            state = {}

            def unf_fwd():
                bgemm(qkv, qkv, S, M=L, N=L, K=dh, lda=3 * H, ldb=3 * H, ldd=Lp, a_mn=False, b_mn=False, batch0=heads,
                      batch1=B, sa=sQ, sb=sQ, sd=sS, b_off=H, alpha=scale)
                state["Pd"] = K.softmax_fwd_(S, mask, B, heads, L, Lp, p, ws.seed, 5)
                bgemm(state["Pd"], qkv, ctxv, M=L, N=dh, K=L, lda=Lp, ldb=3 * H, ldd=H, a_mn=False, b_mn=True,
                      batch0=heads, batch1=B, sa=sS, sb=sQ, sd=sC, b_off=2 * H)

            def unf_bwd():
                Pd = state["Pd"]
                bgemm(dctx, qkv, dP, M=L, N=L, K=dh, lda=H, ldb=3 * H, ldd=Lp, a_mn=False, b_mn=False, batch0=heads,
                      batch1=B, sa=sC, sb=sQ, sd=sS, b_off=2 * H)
                bgemm(Pd, dctx, dqkv, M=L, N=dh, K=L, lda=Lp, ldb=H, ldd=3 * H, a_mn=True, b_mn=True, batch0=heads,
                      batch1=B, sa=sS, sb=sC, sd=sQ, d_off=2 * H)
                K.softmax_bwd_(S, dP, B, heads, L, Lp, p, ws.seed, 5)
                bgemm(dP, qkv, dqkv, M=L, N=dh, K=L, lda=Lp, ldb=3 * H, ldd=3 * H, a_mn=False, b_mn=True, batch0=heads,
                      batch1=B, sa=sS, sb=sQ, sd=sQ, b_off=H, alpha=scale)
                bgemm(dP, qkv, dqkv, M=L, N=dh, K=L, lda=Lp, ldb=3 * H, ldd=3 * H, a_mn=True, b_mn=True, batch0=heads,
                      batch1=B, sa=sS, sb=sQ, sd=sQ, d_off=H, alpha=scale)

            out = torch.empty(B * L, H, device="cuda", dtype=dt)
            dq2 = torch.empty_like(qkv)

            def fu_fwd():
                state["o"], state["lse"] = K.attn_fwd(qkv, mask, B, heads, L, scale, p, ws.seed, 5, Lp=Lp, out=out)

            def fu_bwd():
                K.attn_bwd(qkv, state["o"], dctx, state["lse"], mask, B, heads, L, scale, p, ws.seed, 5, Lp=Lp, out=dq2)

            t_uf, t_ub = timeit(unf_fwd), timeit(unf_bwd)
            t_ff, t_fb = timeit(fu_fwd), timeit(fu_bwd)
            unf_fwd(); unf_bwd(); fu_fwd(); fu_bwd()
            torch.cuda.synchronize()
            err_o = (out.float() - ctxv.float()).abs().max().item() / ctxv.float().abs().max().item()
            err_g = (dq2.float() - dqkv.float()).abs().max().item() / dqkv.float().abs().max().item()
            flops_f = 4.0 * B * heads * L * L * dh
            print(f"B={B} heads={heads} L={L} {str(dt)[6:]:8s} unfused fwd {t_uf:7.1f} us bwd {t_ub:7.1f} us | "
                  f"fused fwd {t_ff:7.1f} us ({flops_f / t_ff / 1e6:6.1f} TF/s) bwd {t_fb:7.1f} us "
                  f"({2.5 * flops_f / t_fb / 1e6:6.1f} TF/s) | rel diff o {err_o:.2e} dqkv {err_g:.2e}", flush=True)


if __name__ == "__main__":
    main()
