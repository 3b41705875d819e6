"""Numerics of every hand-written kernel against a plain PyTorch fp32/fp64 restatement of the same op."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _rel(a, b):
    return (a.double() - b.double()).abs().max().item() / (b.double().abs().max().item() + 1e-12)


@pytest.fixture(scope="module")
def K():
    from hetseq_9cme_b200.ops import kernels
    return kernels


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("majors", [(False, False), (False, True), (True, True), (True, False)])
def test_gemm_layouts(dtype, majors):
    from hetseq_9cme_b200.ops.gemm import gemm
    a_mn, b_mn = majors
    torch.manual_seed(0)
    M, N, K_ = 384, 320, 200 if dtype == torch.float32 else 208
    a = torch.randn((K_, M) if a_mn else (M, K_), device="cuda").to(dtype)
    b = torch.randn((K_, N) if b_mn else (N, K_), device="cuda").to(dtype)
    out = gemm(a, b, a_mn=a_mn, b_mn=b_mn)
    A = a.double().t() if a_mn else a.double()
    Bm = b.double().t() if b_mn else b.double()
    ref = A @ Bm.t()
    assert _rel(out, ref) < (5e-3 if dtype == torch.float32 else 2e-2)


def test_gemm_epilogues_and_accumulate():
    from hetseq_9cme_b200.ops.gemm import gemm
    from hetseq_9cme_b200.ops import native as nat
    torch.manual_seed(1)
    M, N, Kd = 256, 512, 256
    a = torch.randn(M, Kd, device="cuda")
    w = torch.randn(N, Kd, device="cuda") * 0.1
    bias = torch.randn(N, device="cuda")
    aux = torch.randn(M, N, device="cuda")
    base = a.double() @ w.double().t()
    pre = torch.empty(M, N, device="cuda")
    got = gemm(a, w, bias=bias, act=nat.ACT_GELU, preact_out=pre)
    assert _rel(pre, base + bias.double()) < 5e-3
    assert _rel(got, torch.nn.functional.gelu(base + bias.double())) < 5e-3
    assert _rel(gemm(a, w, bias=bias, aux=aux, aux_mode=nat.AUX_ADD), base + bias.double() + aux.double()) < 5e-3
    x = aux.double().requires_grad_(True)
    torch.nn.functional.gelu(x).sum().backward()
    assert _rel(gemm(a, w, aux=aux, aux_mode=nat.AUX_MUL_DGELU), base * x.grad) < 5e-3
    acc = torch.ones(M, N, device="cuda")
    gemm(a, w, out=acc, accumulate=True, splits=2)
    assert _rel(acc, base + 1.0) < 5e-3


def test_batched_gemm_attention_shapes():
    from hetseq_9cme_b200.ops.gemm import bgemm
    torch.manual_seed(2)
    B, h, L, dh = 3, 4, 80, 64
    H = h * dh
    Lp = (L + 7) // 8 * 8
    qkv = torch.randn(B * L, 3 * H, device="cuda")
    S = torch.zeros(B, h, L, Lp, device="cuda")
    bgemm(qkv, qkv, S, M=L, N=L, K=dh, lda=3 * H, ldb=3 * H, ldd=Lp, a_mn=False, b_mn=False, batch0=h, batch1=B,
          sa=(dh, L * 3 * H), sb=(dh, L * 3 * H), sd=(L * Lp, h * L * Lp), b_off=H, alpha=0.125)
    q = qkv.view(B, L, 3, h, dh)[:, :, 0].permute(0, 2, 1, 3).double()
    k = qkv.view(B, L, 3, h, dh)[:, :, 1].permute(0, 2, 1, 3).double()
    v = qkv.view(B, L, 3, h, dh)[:, :, 2].permute(0, 2, 1, 3).double()
    ref = (q @ k.transpose(-1, -2)) * 0.125
    assert _rel(S[..., :L], ref) < 5e-3
    P = torch.softmax(ref, -1).float()
    Pp = torch.zeros(B, h, L, Lp, device="cuda")
    Pp[..., :L] = P
    ctx = torch.empty(B * L, H, device="cuda")
    bgemm(Pp, qkv, ctx, M=L, N=dh, K=L, lda=Lp, ldb=3 * H, ldd=H, a_mn=False, b_mn=True, batch0=h, batch1=B,
          sa=(L * Lp, h * L * Lp), sb=(dh, L * 3 * H), sd=(dh, L * H), b_off=2 * H)
    refc = (P.double() @ v).permute(0, 2, 1, 3).reshape(B * L, H)
    assert _rel(ctx, refc) < 5e-3
    # MN-major A (transposed probabilities) as used by dV
    dctx = torch.randn(B * L, H, device="cuda")
    dqkv = torch.zeros(B * L, 3 * H, device="cuda")
    bgemm(Pp, dctx, dqkv, M=L, N=dh, K=L, lda=Lp, ldb=H, ldd=3 * H, a_mn=True, b_mn=True, batch0=h, batch1=B,
          sa=(L * Lp, h * L * Lp), sb=(dh, L * H), sd=(dh, L * 3 * H), d_off=2 * H)
    dc = dctx.view(B, L, h, dh).permute(0, 2, 1, 3).double()
    refdv = (P.double().transpose(-1, -2) @ dc).permute(0, 2, 1, 3).reshape(B * L, H)
    assert _rel(dqkv[:, 2 * H:], refdv) < 5e-3


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("cols", [128, 768, 1024])
def test_layernorm_fwd_bwd(K, dtype, cols):
    torch.manual_seed(3)
    rows = 515
    x = torch.randn(rows, cols, device="cuda").to(dtype)
    res = torch.randn(rows, cols, device="cuda").to(dtype)
    bias = torch.randn(cols, device="cuda").to(dtype)
    gamma = (torch.rand(cols, device="cuda") + 0.5).to(dtype)
    beta = torch.randn(cols, device="cuda").to(dtype)
    y, mean, rstd = K.ln_fwd(x, gamma, beta, bias=bias, residual=res, eps=1e-12)
    xr = (x.double() + bias.double() + res.double()).requires_grad_(True)
    gr, br = gamma.double().requires_grad_(True), beta.double().requires_grad_(True)
    mu = xr.mean(-1, keepdim=True)
    var = (xr - mu).pow(2).mean(-1, keepdim=True)
    yr = (xr - mu) / torch.sqrt(var + 1e-12) * gr + br
    tol = 1e-4 if dtype == torch.float32 else 2e-2
    assert _rel(y, yr) < tol
    dy = torch.randn(rows, cols, device="cuda").to(dtype)
    yr.backward(dy.double())
    dg = torch.zeros(cols, device="cuda")
    db = torch.zeros(cols, device="cuda")
    dbias = torch.zeros(cols, device="cuda")
    ds, dz = K.ln_bwd(dy, x, gamma, beta, mean, rstd, dg, db, bias=bias, dbias=dbias, residual=res, eps=1e-12)
    assert dz is None
    dg2, db2, dbias2 = torch.zeros(cols, device="cuda"), torch.zeros(cols, device="cuda"), torch.zeros(cols, device="cuda")
    K.ln_bwd(dy, x, gamma, beta, mean, rstd, dg2, db2, bias=bias, dbias=dbias2, residual=res, eps=1e-12, deterministic=True)
    assert _rel(dg2, dg) < 1e-4 and _rel(db2, db) < 1e-4 and _rel(dbias2, dbias) < 1e-4   # two-pass == atomic path
    assert _rel(ds, xr.grad) < tol
    assert _rel(dg, gr.grad) < tol
    assert _rel(db, br.grad) < tol
    assert _rel(dbias, xr.grad.sum(0)) < tol


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("cols,ctas", [(768, 2), (768, 1), (1024, 2), (256, 2)])
def test_layernorm_streaming_kernels_match(K, dtype, cols, ctas):
    """The cp.async-pipelined LayerNorm kernels (b200_ln_config) run the same arithmetic on the same values as the
    direct-load ones: y / mean / rstd / ds / dz bit for bit (pre-dropout on, ragged row count, several rows per
    warp), parameter gradients up to the order of the atomic adds."""
    torch.manual_seed(11)
    rows, p = 4099, 0.1
    ws = K.Workspace.get("cuda")
    ws.set_seed(77)
    x = torch.randn(rows, cols, device="cuda").to(dtype)
    res = torch.randn(rows, cols, device="cuda").to(dtype)
    bias = torch.randn(cols, device="cuda").to(dtype)
    gamma = (torch.rand(cols, device="cuda") + 0.5).to(dtype)
    beta = torch.randn(cols, device="cuda").to(dtype)
    dy = torch.randn(rows, cols, device="cuda").to(dtype)

    def run():
        y, mean, rstd = K.ln_fwd(x, gamma, beta, bias=bias, residual=res, p_pre=p, seed=ws.seed, stream_pre=9)
        dg, db, dbias = (torch.zeros(cols, device="cuda") for _ in range(3))
        ds, dz = K.ln_bwd(dy, x, gamma, beta, mean, rstd, dg, db, bias=bias, dbias=dbias, residual=res, p_pre=p,
                          seed=ws.seed, stream_pre=9, want_dz=True)
        dg2, db2, dbias2 = (torch.zeros(cols, device="cuda") for _ in range(3))
        K.ln_bwd(dy, x, gamma, beta, mean, rstd, dg2, db2, bias=bias, dbias=dbias2, residual=res, p_pre=p,
                 seed=ws.seed, stream_pre=9, want_dz=True, deterministic=True)
        torch.cuda.synchronize()
        return (y, mean, rstd, ds, dz), (dg, db, dbias, dg2, db2, dbias2)

    before = K.ln_config()
    try:
        K.ln_config(0)
        exact0, sums0 = run()
        assert K.ln_config(3, ctas, ctas) == 3
        exact1, sums1 = run()
    finally:
        K.ln_config(before, 2, 2)
    for a, b in zip(exact0, exact1):
        assert torch.equal(a, b)
    for a, b in zip(sums0, sums1):
        assert _rel(b, a) < 1e-5
    assert 0.05 < (exact1[4] == 0).float().mean().item() < 0.15      # dz carries the pre-dropout mask


def test_layernorm_dropout_is_consistent(K):
    """Forward and backward must regenerate the same mask; keep-rate must match p."""
    torch.manual_seed(4)
    rows, cols, p = 2048, 768, 0.1
    ws = K.Workspace.get("cuda")
    ws.set_seed(1234)
    x = torch.randn(rows, cols, device="cuda")
    gamma, beta = torch.ones(cols, device="cuda"), torch.zeros(cols, device="cuda")
    # post-dropout: zeros in y mark dropped elements
    y, mean, rstd = K.ln_fwd(x, gamma, beta, p_post=p, seed=ws.seed, stream_post=5)
    dropped = (y == 0)
    rate = dropped.float().mean().item()
    assert abs(rate - p) < 0.01
    dy = torch.ones_like(x)
    dg, db = torch.zeros(cols, device="cuda"), torch.zeros(cols, device="cuda")
    K.ln_bwd(dy, x, gamma, beta, mean, rstd, dg, db, p_post=p, seed=ws.seed, stream_post=5)
    # d(beta) = sum over rows of the masked dy = kept count * 1/(1-p)
    kept = (~dropped).float().sum(0) / (1 - p)
    assert _rel(db, kept) < 1e-5
    y2, _, _ = K.ln_fwd(x, gamma, beta, p_post=p, seed=ws.seed, stream_post=6)
    assert (y2 == 0).ne(dropped).any()          # a different stream id gives a different mask


def test_softmax_fwd_bwd(K):
    torch.manual_seed(5)
    B, h, L = 2, 3, 100
    Lp = 104
    S = torch.randn(B, h, L, Lp, device="cuda")
    mask = torch.ones(B, L, dtype=torch.int64, device="cuda")
    mask[1, 70:] = 0
    ref_in = S[..., :L].double() + ((1 - mask.double()) * -10000.0)[:, None, None, :]
    ref_in.requires_grad_(True)
    Pr = torch.softmax(ref_in, -1)
    Sg = S.clone()
    K.softmax_fwd_(Sg, mask, B, h, L, Lp)
    assert _rel(Sg[..., :L], Pr) < 1e-5
    assert Sg[..., L:].abs().max().item() == 0
    dP = torch.randn(B, h, L, Lp, device="cuda")
    dP[..., L:] = 0
    Pr.backward(dP[..., :L].double())
    K.softmax_bwd_(Sg, dP, B, h, L, Lp)
    assert _rel(dP[..., :L], ref_in.grad) < 1e-4


def test_cross_entropy_wide(K):
    torch.manual_seed(6)
    rows, V = 96, 30522
    ld = (V + 3) // 4 * 4
    logits = (torch.randn(rows, ld, device="cuda") * 3)[:, :V]        # padded pitch, as produced by gemm()
    labels = torch.randint(0, V, (rows,), device="cuda")
    labels[::5] = -1
    ref_l = logits.double().clone().requires_grad_(True)
    ref = torch.nn.functional.cross_entropy(ref_l, labels, ignore_index=-1)
    lse, row_loss = K.ce_fwd(logits, labels, ignore=-1)
    loss, inv, cnt = K.ce_finalize(row_loss, labels, ignore=-1)
    assert abs(loss.item() - ref.item()) < 1e-4
    ref.backward()
    g = torch.full((1,), 1.0, device="cuda")
    K.ce_bwd_(logits, labels, lse, gout=g, inv_count=inv, ignore=-1)
    assert _rel(logits, ref_l.grad) < 1e-4


def test_embedding_and_scatter(K):
    torch.manual_seed(7)
    B, L, H, V = 4, 32, 128, 1000
    W, P, T = torch.randn(V, H, device="cuda"), torch.randn(64, H, device="cuda"), torch.randn(2, H, device="cuda")
    ids = torch.randint(0, V, (B, L), device="cuda")
    tts = torch.randint(0, 2, (B, L), device="cuda")
    out = K.embed_sum(ids.view(-1), tts.view(-1), W, P, T, L)
    pos = torch.arange(L, device="cuda").repeat(B)
    ref = W[ids.view(-1)] + P[pos] + T[tts.view(-1)]
    assert _rel(out, ref) < 1e-6
    ds = torch.randn(B * L, H, device="cuda")
    gW, gP, gT = torch.zeros_like(W), torch.zeros_like(P), torch.zeros_like(T)
    K.embed_scatter(ids.view(-1), tts.view(-1), ds, gW, gP, gT, L)
    rW = torch.zeros_like(W).index_add_(0, ids.view(-1), ds)
    rP = torch.zeros_like(P).index_add_(0, pos, ds)
    rT = torch.zeros_like(T).index_add_(0, tts.view(-1), ds)
    assert _rel(gW, rW) < 1e-5 and _rel(gP, rP) < 1e-5 and _rel(gT, rT) < 1e-5


def test_compact_gather_scatter(K):
    torch.manual_seed(8)
    rows, H = 3000, 64
    labels = torch.full((rows,), -1, dtype=torch.int64, device="cuda")
    sel = torch.randperm(rows, device="cuda")[:400].sort().values
    labels[sel] = torch.randint(0, 100, (400,), device="cuda")
    idx, lab, cnt = K.compact_rows(labels, 512, ignore=-1)
    assert cnt.tolist() == [400, 400]
    assert torch.equal(idx[:400].long(), sel) and (idx[400:] == -1).all()
    assert torch.equal(lab[:400], labels[sel])
    x = torch.randn(rows, H, device="cuda")
    g = K.gather_rows(x, idx)
    assert torch.equal(g[:400], x[sel]) and g[400:].abs().max().item() == 0
    dx = torch.zeros(rows, H, device="cuda")
    K.scatter_rows(g, idx, dx)
    assert torch.equal(dx[sel], x[sel])


def test_small_heads(K):
    torch.manual_seed(9)
    rows, Kd, N = 700, 256, 9
    x = torch.randn(rows, Kd, device="cuda")
    w = torch.randn(N, Kd, device="cuda") * 0.1
    b = torch.randn(N, device="cuda")
    labels = torch.randint(0, N, (rows,), device="cuda")
    labels[::7] = -100
    active = torch.ones(rows, dtype=torch.int64, device="cuda")
    active[::3] = 0
    xr, wr, br = x.double().requires_grad_(True), w.double().requires_grad_(True), b.double().requires_grad_(True)
    lr_ = xr @ wr.t() + br
    eff = torch.where(active == 1, labels, torch.full_like(labels, -100))
    ref = torch.nn.functional.cross_entropy(lr_, eff, ignore_index=-100)
    ref.backward()
    logits = K.head_fwd(x, w, b)
    assert _rel(logits, lr_) < 1e-5
    rl, effk = K.head_ce(logits, labels, active, ignore=-100)
    loss, inv, _ = K.ce_finalize(rl, effk, ignore=-100)
    assert abs(loss.item() - ref.item()) < 1e-5
    dl = K.head_ce_bwd(logits, effk, gout=None, inv_count=inv, ignore=-100)
    dx = torch.empty_like(x)
    dw, db = torch.zeros_like(w), torch.zeros_like(b)
    K.head_bwd(dl, x, w, dx=dx, dw=dw, db=db)
    assert _rel(dx, xr.grad) < 1e-4 and _rel(dw, wr.grad) < 1e-4 and _rel(db, br.grad) < 1e-4


def test_colsum_actbwd_dropout(K):
    torch.manual_seed(10)
    x = torch.randn(4096, 768, device="cuda")
    out = torch.ones(768, device="cuda")
    K.colsum_(x, out)
    assert _rel(out, x.double().sum(0) + 1) < 1e-5
    u = torch.randn(512, 768, device="cuda")
    ur = u.double().requires_grad_(True)
    torch.nn.functional.gelu(ur).sum().backward()
    dy = torch.randn_like(u)
    assert _rel(K.act_bwd(dy, u, 1), dy.double() * ur.grad) < 1e-5
    ws = K.Workspace.get("cuda")
    ws.set_seed(77)
    y = K.dropout(x, 0.25, ws.seed, 3)
    keep = (y != 0)
    assert abs(keep.float().mean().item() - 0.75) < 0.01
    assert _rel(y[keep], x[keep] / 0.75) < 1e-6
    assert torch.equal(K.dropout(x, 0.25, ws.seed, 3), y)


def test_sumsq_and_adam_match_reference_math(K):
    torch.manual_seed(11)
    n = 1_000_003
    npad = (n + 63) // 64 * 64
    p = torch.randn(npad, device="cuda")
    g = torch.randn(npad, device="cuda") * 0.01
    m = torch.zeros(npad, device="cuda")
    v = torch.zeros(npad, device="cuda")
    ws = K.Workspace.get("cuda")
    K.sumsq(g, ws.norm_sq)
    assert abs(ws.norm_sq.item() - g.double().pow(2).sum().item()) / g.double().pow(2).sum().item() < 1e-5
    # reference Adam math (hetseq/optim.py:206-229) in float64
    lr, b1, b2, eps, wd = 1e-3, 0.9, 0.999, 1e-8, 0.01
    mult, ss, max_norm = 8.0, 128.0, 0.05
    pr, gr, mr, vr = p.double(), g.double(), m.double(), v.double()
    c = mult / ss
    gnorm = gr.norm().item() * c
    coef = max_norm / (gnorm + 1e-6) if gnorm > max_norm else 1.0
    for step in (1, 2):
        gs = gr * c * coef
        mr = b1 * mr + (1 - b1) * gs
        vr = b2 * vr + (1 - b2) * gs * gs
        denom = vr.sqrt() + eps
        step_size = lr * math.sqrt(1 - b2 ** step) / (1 - b1 ** step)
        pr = pr * (1 - wd * lr) - step_size * mr / denom
    sample = torch.tensor([ss], dtype=torch.float64, device="cuda")
    stats = torch.zeros(4, device="cuda")
    for step in (1, 2):
        K.sumsq(g, ws.norm_sq)
        K.adam_flat(p, g, m, v, npad, lr=lr, beta1=b1, beta2=b2, eps=eps, weight_decay=wd, step=step, grad_scale_mult=mult,
                    max_norm=max_norm, norm_sq=ws.norm_sq, sample_size=sample, stats_out=stats, zero_grad=False)
    assert _rel(p, pr) < 1e-5
    assert abs(stats[0].item() - gnorm) / gnorm < 1e-4 and stats[2].item() == 1.0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("L", [128, 72])
def test_softmax_epilogues_match_separate_kernels(K, dtype, L):
    """Score GEMM with the softmax epilogue == plain score GEMM followed by the softmax kernel (fwd and bwd),
    including identical dropout masks."""
    from hetseq_9cme_b200.ops.gemm import bgemm
    torch.manual_seed(12)
    B, h, dh, p = 2, 3, 64, 0.1
    H = h * dh
    Lp = (L + 7) // 8 * 8
    ws = K.Workspace.get("cuda")
    ws.set_seed(4321)
    qkv = torch.randn(B * L, 3 * H, device="cuda").to(dtype)
    mask = torch.ones(B, L, dtype=torch.int64, device="cuda")
    mask[1, L - 11:] = 0
    common = dict(M=L, N=L, K=dh, lda=3 * H, ldb=3 * H, ldd=Lp, a_mn=False, b_mn=False, batch0=h, batch1=B,
                  sa=(dh, L * 3 * H), sb=(dh, L * 3 * H), sd=(L * Lp, h * L * Lp), b_off=H, alpha=0.125)
    S = torch.zeros(B, h, L, Lp, device="cuda", dtype=dtype)
    bgemm(qkv, qkv, S, **common)
    Pd_ref = K.softmax_fwd_(S, mask, B, h, L, Lp, p, ws.seed, 77)
    P = torch.zeros_like(S); Pd = torch.zeros_like(S)
    bgemm(qkv, qkv, P, epi_mode=1, attn_mask=mask, p_drop=p, seed=ws.seed, stream_id=77, d2=Pd, **common)
    tol = 1e-5 if dtype == torch.float32 else 1e-2
    assert (P[..., :L].float() - S[..., :L].float()).abs().max().item() < tol
    assert (Pd[..., :L].float() - Pd_ref[..., :L].float()).abs().max().item() < tol * 2
    assert ((Pd[..., :L] == 0) == (Pd_ref[..., :L] == 0)).float().mean().item() > 0.999      # same dropout mask
    # backward: dS from the epilogue of dP = dctx V^T
    dctx = torch.randn(B * L, H, device="cuda").to(dtype)
    bw = dict(M=L, N=L, K=dh, lda=H, ldb=3 * H, ldd=Lp, a_mn=False, b_mn=False, batch0=h, batch1=B,
              sa=(dh, L * H), sb=(dh, L * 3 * H), sd=(L * Lp, h * L * Lp), b_off=2 * H)
    dP = torch.zeros_like(S)
    bgemm(dctx, qkv, dP, **bw)
    K.softmax_bwd_(S, dP, B, h, L, Lp, p, ws.seed, 77)
    dS = torch.zeros_like(S)
    bgemm(dctx, qkv, dS, epi_mode=2, aux=S, p_drop=p, seed=ws.seed, stream_id=77, **bw)
    scale = dP[..., :L].float().abs().max().item()
    assert (dS[..., :L].float() - dP[..., :L].float()).abs().max().item() < (1e-4 if dtype == torch.float32 else 3e-2) * scale


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("L,p", [(128, 0.0), (128, 0.1), (72, 0.1), (320, 0.1)])
def test_fused_attention_matches_reference(K, dtype, L, p):
    """Flash-style tcgen05 attention (forward + dQ/dK/dV) against fp64 PyTorch attention that applies the very same
    dropout mask (read back from the stand-alone softmax kernel, which shares the Philox indexing)."""
    torch.manual_seed(21)
    B, h, dh = 2, 3, 64
    H = h * dh
    Lp = (L + 3) // 4 * 4
    ws = K.Workspace.get("cuda")
    ws.set_seed(977)
    qkv = (torch.randn(B * L, 3 * H, device="cuda") * 0.7).to(dtype)
    dout = torch.randn(B * L, H, device="cuda").to(dtype)
    mask = torch.ones(B, L, dtype=torch.int64, device="cuda")
    mask[1, L - 13:] = 0
    scale = 1.0 / math.sqrt(dh)

    keep = torch.ones(B, h, L, L, device="cuda", dtype=torch.float64)
    if p > 0:
        probe = torch.ones(B, h, L, Lp, device="cuda")
        keep = (K.softmax_fwd_(probe, None, B, h, L, Lp, p, ws.seed, 33)[..., :L] != 0).double() / (1 - p)

    x = qkv.double().view(B, L, 3, h, dh).requires_grad_(True)
    q, k, v = (x[:, :, i].permute(0, 2, 1, 3) for i in range(3))
    s = q @ k.transpose(-1, -2) * scale + ((1 - mask.double()) * -10000.0)[:, None, None, :]
    o_ref = ((torch.softmax(s, -1) * keep) @ v).permute(0, 2, 1, 3).reshape(B * L, H)
    o_ref.backward(dout.double())
    dqkv_ref = x.grad.reshape(B * L, 3 * H)

    o, lse = K.attn_fwd(qkv, mask, B, h, L, scale, p, ws.seed, 33, Lp=Lp)
    dqkv = K.attn_bwd(qkv, o, dout, lse, mask, B, h, L, scale, p, ws.seed, 33, Lp=Lp)
    torch.cuda.synchronize()
    tol = 2e-3 if dtype == torch.float32 else 2e-2       # tf32 / bf16 operand rounding
    assert _rel(o, o_ref) < tol
    lse_ref = torch.logsumexp(s, -1).reshape(-1)
    assert (lse.double() - lse_ref).abs().max().item() < (5e-3 if dtype == torch.float32 else 5e-2)
    for i, name in enumerate("qkv"):
        got = dqkv.view(B * L, 3, H)[:, i]
        want = dqkv_ref.view(B * L, 3, H)[:, i]
        assert _rel(got, want) < tol * 2, name


class _Guarded:
    """Output tensor carved out of a larger buffer whose surroundings hold a sentinel: a cheap stand-in for
    ``compute-sanitizer --tool memcheck`` (closed on the development pool) that catches out-of-bounds WRITES."""

    PAD = 4096

    def __init__(self, shape, dtype, pitch=None):
        rows = shape[0] if len(shape) > 1 else 1
        cols = shape[-1]
        self.pitch = pitch or cols
        n = rows * self.pitch
        self.buf = torch.full((n + 2 * self.PAD,), 7.0, device="cuda").to(dtype)
        body = self.buf[self.PAD:self.PAD + n]
        self.view = body.view(rows, self.pitch)[:, :cols] if len(shape) > 1 else body[:cols]
        self.rows, self.cols = rows, cols

    def check(self, what):
        b = self.buf.float()
        assert bool((b[:self.PAD] == 7.0).all()) and bool((b[-self.PAD:] == 7.0).all()), what + ": wrote outside the buffer"
        if self.pitch != self.cols:
            gaps = b[self.PAD:-self.PAD].view(self.rows, self.pitch)[:, self.cols:]
            assert bool((gaps == 7.0).all()), what + ": wrote into the pitch padding"
        assert bool((self.view.float() != 7.0).any()), what + ": wrote nothing"


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_kernels_respect_output_bounds(K, dtype):
    from hetseq_9cme_b200.ops.gemm import gemm
    torch.manual_seed(3)
    ws = K.Workspace.get("cuda")
    ws.set_seed(11)
    # GEMM with ragged M / N (TMA store clipping) into a pitched output
    for (M, N, Kd) in [(100, 72, 64), (257, 200, 96), (4096, 24, 128)]:
        a = torch.randn(M, Kd, device="cuda").to(dtype)
        b = torch.randn(N, Kd, device="cuda").to(dtype)
        g = _Guarded((M, N), dtype, pitch=(N + 7) // 8 * 8 + 8)
        gemm(a, b, out=g.view)
        torch.cuda.synchronize()
        g.check("gemm %dx%dx%d" % (M, N, Kd))
        assert _rel(g.view, a.float() @ b.float().t()) < (2e-3 if dtype == torch.float32 else 2e-2)
    # fused attention, ragged L (TMA store clipping of rows >= L, head-strided writes into [B*L, 3H])
    for L in (72, 128, 200):
        B, h = 2, 3
        H = h * 64
        qkv = torch.randn(B * L, 3 * H, device="cuda").to(dtype)
        dout = torch.randn(B * L, H, device="cuda").to(dtype)
        go = _Guarded((B * L, H), dtype)
        o, lse = K.attn_fwd(qkv, None, B, h, L, 0.125, 0.1, ws.seed, 3, out=go.view)
        gd = _Guarded((B * L, 3 * H), dtype)
        K.attn_bwd(qkv, o, dout, lse, None, B, h, L, 0.125, 0.1, ws.seed, 3, out=gd.view)
        torch.cuda.synchronize()
        go.check("attn_fwd L=%d" % L)
        gd.check("attn_bwd L=%d" % L)
        assert bool(torch.isfinite(gd.view.float()).all()) and bool((gd.view.float() != 7.0).all())
    # batch assembly
    if dtype == torch.float32:
        B, L, P = 5, 96, 12
        raw = torch.randint(1, 90, (B, 3 * L + 2 * P + 1), dtype=torch.int32, device="cuda")
        outs = [torch.full((B * L + 64,), 7, dtype=torch.int64, device="cuda") for _ in range(4)]
        nsl = torch.full((B + 64,), 7, dtype=torch.int64, device="cuda")
        K.assemble_bert_batch(raw, L, P, out=[t[:B * L].view(B, L) for t in outs] + [nsl[:B]])
        torch.cuda.synchronize()
        assert all(bool((t[B * L:] == 7).all()) for t in outs) and bool((nsl[B:] == 7).all())


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("majors", [(False, False), (False, True), (True, True), (True, False)])
def test_gemm_cta_pair_matches_single_cta(dtype, majors, monkeypatch):
    """The cta_group::2 kernel (two CTAs share each tcgen05.mma, each loads half of B) against the single-CTA kernel and
    fp64, on shapes with ragged M / N, a bias + GELU epilogue and fp32 accumulation into an existing output."""
    from hetseq_9cme_b200.ops import native as N
    from hetseq_9cme_b200.ops.gemm import gemm
    a_mn, b_mn = majors
    torch.manual_seed(4)
    for (M, Nn, Kd, bn) in [(512, 256, 128, 256), (1000, 200, 96, 128), (4096, 768, 768, 0), (304, 520, 64, 256)]:
        a = torch.randn((Kd, M) if a_mn else (M, Kd), device="cuda").to(dtype)
        b = torch.randn((Kd, Nn) if b_mn else (Nn, Kd), device="cuda").to(dtype)
        bias = torch.randn(Nn, device="cuda").to(dtype)
        outs = {}
        for flag in ("0", "1"):
            monkeypatch.setenv("B200_GEMM_CTA2", flag)
            outs[flag] = gemm(a, b, a_mn=a_mn, b_mn=b_mn, bias=bias, act=N.ACT_GELU, bn=bn)
        torch.cuda.synchronize()
        A = (a.t() if a_mn else a).double()
        B = (b.t() if b_mn else b).double()
        ref = torch.nn.functional.gelu(A @ B.t() + bias.double())
        tol = 2e-3 if dtype == torch.float32 else 2e-2
        assert _rel(outs["1"], ref) < tol, (M, Nn, Kd)
        assert _rel(outs["1"], outs["0"].double()) < tol / 4
    # accumulate (weight-gradient) mode with split-K
    monkeypatch.setenv("B200_GEMM_CTA2", "1")
    x = torch.randn(4096, 768, device="cuda").to(dtype)
    dy = torch.randn(4096, 512, device="cuda").to(dtype)
    acc = torch.ones(512, 768, device="cuda")
    gemm(dy, x, a_mn=True, b_mn=True, out=acc, out_dtype=torch.float32, accumulate=True)
    torch.cuda.synchronize()
    assert _rel(acc, 1.0 + dy.double().t() @ x.double()) < (2e-3 if dtype == torch.float32 else 1e-2)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_gemm_gelu_with_saved_derivative(dtype):
    """FFN-1 epilogue that emits gelu(x) and gelu'(x) together (one exponential, rational erf) and the backward epilogue
    that multiplies by the stored derivative, against torch's erf-GELU and autograd."""
    from hetseq_9cme_b200.ops import native as N
    from hetseq_9cme_b200.ops.gemm import gemm
    torch.manual_seed(8)
    M, H, I = 512, 256, 1024
    x = torch.randn(M, H, device="cuda").to(dtype)
    w = (torch.randn(I, H, device="cuda") * 0.1).to(dtype)
    b = torch.randn(I, device="cuda").to(dtype)
    dact = torch.empty(M, I, device="cuda", dtype=dtype)
    y = gemm(x, w, bias=b, act=N.ACT_GELU_SAVE_GRAD, preact_out=dact)
    pre = (x.double() @ w.double().t() + b.double()).requires_grad_(True)
    ref = torch.nn.functional.gelu(pre)
    g = torch.randn(M, I, device="cuda", dtype=torch.float64)
    ref.backward(g)
    tol = 2e-3 if dtype == torch.float32 else 2e-2
    assert _rel(y, ref.detach()) < tol
    # gelu'(pre): |gelu''| <= 1.13, so the operand rounding of `pre` (TF32: ~2e-3 here) carries over one to one
    assert (dact.double() - pre.grad / g).abs().max().item() < 3 * tol
    # backward: (dz W2) * gelu'  with W2 [H, I]
    dz = torch.randn(M, H, device="cuda").to(dtype)
    w2 = (torch.randn(H, I, device="cuda") * 0.1).to(dtype)
    du = gemm(dz, w2, b_mn=True, aux=dact, aux_mode=N.AUX_MUL)
    want = (dz.double() @ w2.double()) * dact.double()
    assert _rel(du, want) < tol
    # the fast erf form itself: absolute error of gelu against torch over a wide range
    if dtype == torch.float32:
        xs = torch.linspace(-8, 8, 1024 * 64, device="cuda").view(1024, 64)
        eye = torch.eye(64, device="cuda")
        scratch = torch.empty(1024, 64, device="cuda")
        got = gemm(xs, eye, act=N.ACT_GELU_SAVE_GRAD, preact_out=scratch)
        assert (got - torch.nn.functional.gelu(xs.double()).float()).abs().max().item() < 4e-3     # tf32 rounding of xs dominates


def test_strict_fp32_gemm_is_fp32_accurate():
    """--precision fp32: every GEMM as three TF32 products of split operands.  Against an fp64 product the error must
    be at the fp32 level (~1e-6 of |A||B|), two to three orders below the single TF32 product, for plain, epilogue,
    MN-major and accumulating calls."""
    from hetseq_9cme_b200.ops import gemm as G
    from hetseq_9cme_b200.ops import native as N
    torch.manual_seed(3)
    M, Nn, Kk = 384, 320, 1024
    a = torch.randn(M, Kk, device="cuda")
    b = torch.randn(Nn, Kk, device="cuda")
    bias = torch.randn(Nn, device="cuda")
    ref = a.double() @ b.double().t()
    scale = (a.double().abs() @ b.double().abs().t()).max().item()
    tf32 = G.gemm(a, b)
    G.set_strict_fp32(True)
    try:
        strict = G.gemm(a, b)
        e_strict = (strict.double() - ref).abs().max().item() / scale
        e_tf32 = (tf32.double() - ref).abs().max().item() / scale
        assert e_strict < 2e-6 and e_strict < e_tf32 / 50, (e_strict, e_tf32)
        # bias + GELU with the saved derivative, tanh, aux add
        u = torch.empty(M, Nn, device="cuda")
        y = G.gemm(a, b, bias=bias, act=N.ACT_GELU_SAVE_GRAD, preact_out=u, alpha=0.05)
        z = (ref * 0.05 + bias.double())
        assert (y.double() - torch.nn.functional.gelu(z)).abs().max().item() < 1e-5
        cdf = 0.5 * (1 + torch.erf(z / math.sqrt(2)))
        assert (u.double() - (cdf + z * torch.exp(-0.5 * z * z) / math.sqrt(2 * math.pi))).abs().max().item() < 1e-5
        aux = torch.randn(M, Nn, device="cuda")
        w = G.gemm(a, b, aux=aux, aux_mode=N.AUX_ADD)
        assert (w.double() - (ref + aux.double())).abs().max().item() / scale < 2e-6
        # weight-gradient form: A and B MN-major, accumulate into an fp32 buffer
        g = torch.randn(Kk, M, device="cuda"); x = torch.randn(Kk, Nn, device="cuda")
        acc = torch.ones(M, Nn, device="cuda")
        G.gemm(g, x, a_mn=True, b_mn=True, out=acc, accumulate=True)
        want = 1.0 + g.double().t() @ x.double()
        s2 = (g.double().abs().t() @ x.double().abs()).max().item()
        assert (acc.double() - want).abs().max().item() / s2 < 2e-6
    finally:
        G.set_strict_fp32(False)
