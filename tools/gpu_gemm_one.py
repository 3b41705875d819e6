"""Run one GEMM shape a few times (ncu target).  usage: gpu_gemm_one.py M N K [tf32|bf16] [kk|kmn|mnmn] [gelu]

``gelu`` adds the FFN-1 epilogue (bias, GELU, second output with gelu'), i.e. the 8-epilogue-warp CTA-pair variant."""
import sys

import torch

sys.path.insert(0, ".")
from hetseq_9cme_b200.ops.gemm import gemm            # noqa: E402

M, N, K = (int(x) for x in sys.argv[1:4])
dt = torch.bfloat16 if len(sys.argv) > 4 and sys.argv[4] == "bf16" else torch.float32
layout = sys.argv[5] if len(sys.argv) > 5 else "kk"
a_mn, b_mn = layout in ("mnmn",), layout in ("kmn", "mnmn")
a = torch.randn((K, M) if a_mn else (M, K), device="cuda").to(dt)
b = torch.randn((K, N) if b_mn else (N, K), device="cuda").to(dt)
out = torch.empty(M, N, device="cuda", dtype=dt)
epi = {}
if len(sys.argv) > 6 and sys.argv[6] == "gelu":
    from hetseq_9cme_b200.ops import native as Nn
    epi = dict(bias=torch.randn(N, device="cuda").to(dt), act=Nn.ACT_GELU_SAVE_GRAD, preact_out=torch.empty(M, N, device="cuda", dtype=dt))
for _ in range(5):
    gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, **epi)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    gemm(a, b, a_mn=a_mn, b_mn=b_mn, out=out, **epi)
e1.record()
torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 20 * 1e3
print("M=%d N=%d K=%d %s %s: %.1f us  %.0f TFLOP/s" % (M, N, K, str(dt)[6:], layout, us, 2.0 * M * N * K / us / 1e6))
