"""Cost of the GEMM epilogue variants on the FFN shapes (device time inside a CUDA graph)."""
import sys

import torch

sys.path.insert(0, ".")
from hetseq_9cme_b200.ops import native as N          # noqa: E402
from hetseq_9cme_b200.ops.gemm import gemm            # noqa: E402
from tools.gpu_sweep_wgrad import timeit              # noqa: E402


def main():
    for dt in (torch.float32, torch.bfloat16):
        T, H, I = 4096, 768, 3072
        x = torch.randn(T, H, device="cuda").to(dt)
        w1 = torch.randn(I, H, device="cuda").to(dt)
        b1 = torch.randn(I, device="cuda").to(dt)
        out = torch.empty(T, I, device="cuda", dtype=dt)
        d2 = torch.empty(T, I, device="cuda", dtype=dt)
        dz = torch.randn(T, H, device="cuda").to(dt)
        w2 = torch.randn(H, I, device="cuda").to(dt)
        res = torch.randn(T, I, device="cuda").to(dt)
        cases = [
            ("fwd plain", lambda: gemm(x, w1, out=out)),
            ("fwd +bias", lambda: gemm(x, w1, bias=b1, out=out)),
            ("fwd +bias+gelu", lambda: gemm(x, w1, bias=b1, act=N.ACT_GELU, out=out)),
            ("fwd +bias+D2(pre)", lambda: gemm(x, w1, bias=b1, out=out, preact_out=d2)),
            ("fwd +bias+gelu+D2(pre)", lambda: gemm(x, w1, bias=b1, act=N.ACT_GELU, out=out, preact_out=d2)),
            ("fwd +bias+gelu+D2(dgelu)", lambda: gemm(x, w1, bias=b1, act=N.ACT_GELU_SAVE_GRAD, out=out, preact_out=d2)),
            ("dgrad plain", lambda: gemm(dz, w2, b_mn=True, out=out)),
            ("dgrad *aux", lambda: gemm(dz, w2, b_mn=True, out=out, aux=d2, aux_mode=N.AUX_MUL)),
            ("dgrad +aux", lambda: gemm(dz, w2, b_mn=True, out=out, aux=res, aux_mode=N.AUX_ADD)),
            ("dgrad *dgelu(aux)", lambda: gemm(dz, w2, b_mn=True, out=out, aux=d2, aux_mode=N.AUX_MUL_DGELU)),
        ]
        for name, fn in cases:
            print("%-9s %-26s %6.1f us" % (str(dt)[6:], name, timeit(fn)), flush=True)


if __name__ == "__main__":
    main()
