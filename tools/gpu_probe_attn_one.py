"""Run the fused attention forward + backward a few times on one shape (ncu target)."""
import math
import sys

import torch

sys.path.insert(0, ".")
from hetseq_9cme_b200.ops import kernels as K          # noqa: E402

dt = torch.bfloat16 if len(sys.argv) > 1 and sys.argv[1] == "bf16" else torch.float32
B, heads, L, dh = 32, 12, int(sys.argv[2]) if len(sys.argv) > 2 else 128, 64
H = heads * dh
ws = K.Workspace.get("cuda")
ws.set_seed(1)
qkv = torch.randn(B * L, 3 * H, device="cuda").to(dt)
dctx = torch.randn(B * L, H, device="cuda").to(dt)
mask = torch.ones(B, L, dtype=torch.int64, device="cuda")
for _ in range(3):
    o, lse = K.attn_fwd(qkv, mask, B, heads, L, 1 / math.sqrt(dh), 0.1, ws.seed, 5)
    dq = K.attn_bwd(qkv, o, dctx, lse, mask, B, heads, L, 1 / math.sqrt(dh), 0.1, ws.seed, 5)
torch.cuda.synchronize()
print("ok", float(o.float().abs().mean()), float(dq.float().abs().mean()))
