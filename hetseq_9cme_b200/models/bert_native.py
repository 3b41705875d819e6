"""Hand-scheduled sm_100a execution of the BERT blocks.

Each block (embeddings, one encoder layer, the pre-training heads + losses, the token-classification head +
loss) is ONE ``torch.autograd.Function`` whose forward and backward are explicit sequences of the kernels in
``csrc/`` — tcgen05 GEMMs with fused bias / GELU / residual / dGELU epilogues, fused dropout-residual-LayerNorm,
softmax with regenerated dropout masks, fused cross-entropy.  Weight and bias gradients are written straight
into the fp32 gradient arena (``param.main_grad``) by the kernels (TMA reduce-add for the GEMMs), so autograd
only ever carries activation gradients between blocks.

Numerical contract = the reference module it replaces (cited per function); layouts are ours:
activations are 2-D ``[B*L, H]`` row-major throughout, Q/K/V live in one ``[B*L, 3H]`` tensor.
"""
from __future__ import annotations

import math
from typing import List, Optional

import torch

from ..ops import kernels as K
from ..ops import native as N
from ..ops.gemm import bgemm, gemm


class GradSink:
    """Receives "these parameters' gradients for this backward pass are final" notifications; the
    data-parallel reducer subclasses it to launch bucket reductions while backward is still running."""

    def ready(self, params) -> None:  # pragma: no cover - default is a no-op
        pass

    def need(self, params) -> None:  # pragma: no cover - default is a no-op
        """Forward is about to read these parameters (a pipelined optimizer step may still be writing later ones)."""


_NULL_SINK = GradSink()


class NativeCtx:
    """Per-forward execution context shared by the blocks of one model call."""

    def __init__(self, device, training: bool, sink: Optional[GradSink] = None, checkpoint_activations: bool = False):
        self.ws = K.Workspace.get(device)
        self.side = K.SideWork.get(device)
        self.training = training
        self.sink = sink or _NULL_SINK
        self.checkpoint_activations = checkpoint_activations

    @property
    def seed(self):
        return self.ws.seed


def _mg(p) -> torch.Tensor:
    g = getattr(p, "main_grad", None)
    if g is None:
        raise RuntimeError("native BERT path needs parameters bound to a ParamArena (missing main_grad)")
    return g


# ---------------------------------------------------------------------------------------------------
# embeddings  (ref: hetseq/bert_modeling.py:306-320)
# ---------------------------------------------------------------------------------------------------
class EmbeddingsFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, anchor, nctx: NativeCtx, emb, input_ids, token_type_ids, L: int, sid: int):
        ids = input_ids.reshape(-1).contiguous()
        tts = token_type_ids.reshape(-1).contiguous() if token_type_ids is not None else None
        W, P, T = emb.word_embeddings.weight, emb.position_embeddings.weight, emb.token_type_embeddings.weight
        nctx.sink.need([W, P, T, emb.LayerNorm.weight, emb.LayerNorm.bias])
        s = K.embed_sum(ids, tts, W, P, T, L)
        p = emb.dropout.p if nctx.training else 0.0
        y, mean, rstd = K.ln_fwd(s, emb.LayerNorm.weight, emb.LayerNorm.bias, eps=emb.LayerNorm.variance_epsilon,
                                 p_post=p, seed=nctx.seed, stream_post=sid)
        ctx.nctx, ctx.emb, ctx.L, ctx.sid, ctx.p = nctx, emb, L, sid, p
        ctx.save_for_backward(ids, tts if tts is not None else ids.new_empty(0), s, mean, rstd)
        ctx.has_tt = tts is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        ids, tts, s, mean, rstd = ctx.saved_tensors
        emb, nctx = ctx.emb, ctx.nctx
        dy = dy.contiguous()
        ds, _ = K.ln_bwd(dy, s, emb.LayerNorm.weight, emb.LayerNorm.bias, mean, rstd, _mg(emb.LayerNorm.weight),
                         _mg(emb.LayerNorm.bias), eps=emb.LayerNorm.variance_epsilon, p_post=ctx.p, seed=nctx.seed,
                         stream_post=ctx.sid, ws=nctx.ws)
        K.embed_scatter(ids, tts if ctx.has_tt else None, ds, _mg(emb.word_embeddings.weight),
                        _mg(emb.position_embeddings.weight), _mg(emb.token_type_embeddings.weight), ctx.L)
        nctx.sink.ready([emb.position_embeddings.weight, emb.token_type_embeddings.weight, emb.LayerNorm.weight,
                         emb.LayerNorm.bias, emb.word_embeddings.weight])
        return (None,) * 7


# ---------------------------------------------------------------------------------------------------
# one encoder layer  (ref: hetseq/bert_modeling.py:323-441)
# ---------------------------------------------------------------------------------------------------
class LayerPack:
    """Views of one BertLayer's parameters in the layout the kernels want (built once per arena)."""

    def __init__(self, layer, arena):
        att, out = layer.attention.self, layer.attention.output
        self.params = list(layer.parameters())
        qkv_w = [att.query.weight, att.key.weight, att.value.weight]
        qkv_b = [att.query.bias, att.key.bias, att.value.bias]
        self.Wqkv, self.gWqkv = arena.packed(qkv_w, "data"), arena.packed(qkv_w, "grad")
        self.bqkv, self.gbqkv = arena.packed(qkv_b, "data"), arena.packed(qkv_b, "grad")
        self.Wo, self.bo = out.dense.weight, out.dense.bias
        self.g1, self.be1 = out.LayerNorm.weight, out.LayerNorm.bias
        self.W1, self.b1 = layer.intermediate.dense_act.weight, layer.intermediate.dense_act.bias
        self.W2, self.b2 = layer.output.dense.weight, layer.output.dense.bias
        self.g2, self.be2 = layer.output.LayerNorm.weight, layer.output.LayerNorm.bias
        self.eps = out.LayerNorm.variance_epsilon
        self.heads = att.num_attention_heads
        self.dh = att.attention_head_size
        self.p_attn = att.dropout.p
        self.p_hidden = out.dropout.p


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


def _fused_softmax(L: int) -> bool:
    """Softmax inside the GEMM epilogue needs a whole score row in one 128 x BN tile (BN <= 256)."""
    import os
    # measured on B200 (profiles/README.md): at L = 128 the four epilogue warps become the bottleneck and the fused
    # form is ~6 % slower per step than the separate full-chip softmax kernels, so it is opt-in for now
    from ..ops.gemm import strict_fp32
    return L <= 256 and bool(os.environ.get("B200_FUSED_SOFTMAX")) and not strict_fp32()


def _fused_attention(L: int, dh: int) -> bool:
    """Flash-style attention kernels (csrc/attention_sm100.cu): no [B, heads, L, L] tensor is written at all.

    Measured on B200 (tools/gpu_probe_attn.py, profiles/README.md): for L <= 128 the fused forward + single-kernel
    backward take 72 us per BERT-base layer against 103 us for GEMM + softmax + GEMM (tf32; 57 vs 76 us in bf16).
    For longer sequences the two-kernel backward is still slower than the unfused GEMMs, so it is opt-in there
    (B200_FUSED_ATTN=1; it saves the 2 * B * heads * L * L activation tensors).  B200_FUSED_ATTN=0 disables it."""
    import os
    flag = os.environ.get("B200_FUSED_ATTN", "")
    from ..ops.gemm import strict_fp32
    if dh != 64 or flag == "0" or strict_fp32():     # the fused kernels multiply in TF32: strict fp32 takes the GEMM path
        return False
    return L <= 128 or flag == "1"


def _attention_unfused(qkv, nctx, attn_mask, B, L, Lp, H, heads, dh, sid, p_attn):
    """Score GEMM -> softmax (+mask, +dropout) -> context GEMM; returns (ctx, P, Pdrop)."""
    dev, dt = qkv.device, qkv.dtype
    S = torch.empty((B, heads, L, Lp), dtype=dt, device=dev)
    if _fused_softmax(L):
        # softmax (+ mask, + dropout copy) runs in the epilogue of the score GEMM: the scores never reach memory
        Pd = torch.empty_like(S) if p_attn > 0 else S
        bgemm(qkv, qkv, S, M=L, N=L, K=dh, lda=3 * H, ldb=3 * H, ldd=Lp, a_mn=False, b_mn=False, batch0=heads, batch1=B,
              sa=(dh, L * 3 * H), sb=(dh, L * 3 * H), sd=(L * Lp, heads * L * Lp), b_off=H, alpha=1.0 / math.sqrt(dh),
              epi_mode=1, attn_mask=attn_mask, p_drop=p_attn, seed=nctx.seed, stream_id=sid + 1,
              d2=Pd if p_attn > 0 else None)
    else:
        bgemm(qkv, qkv, S, M=L, N=L, K=dh, lda=3 * H, ldb=3 * H, ldd=Lp, a_mn=False, b_mn=False, batch0=heads, batch1=B,
              sa=(dh, L * 3 * H), sb=(dh, L * 3 * H), sd=(L * Lp, heads * L * Lp), b_off=H, alpha=1.0 / math.sqrt(dh))
        Pd = K.softmax_fwd_(S, attn_mask, B, heads, L, Lp, p_attn, nctx.seed, sid + 1)       # S now holds P
    ctxv = torch.empty((B * L, H), dtype=dt, device=dev)
    bgemm(Pd, qkv, ctxv, M=L, N=dh, K=L, lda=Lp, ldb=3 * H, ldd=H, a_mn=False, b_mn=True, batch0=heads, batch1=B,
          sa=(L * Lp, heads * L * Lp), sb=(dh, L * 3 * H), sd=(dh, L * H), b_off=2 * H)
    return ctxv, S, Pd


def _layer_forward(x, nctx, lp, attn_mask, B, L, sid, p_attn, p_hid):
    """Forward kernels of one encoder layer; returns (output, tensors backward needs)."""
    H = x.shape[1]
    heads, dh = lp.heads, lp.dh
    Lp = _pad8(L)
    dev, dt = x.device, x.dtype
    nctx.sink.need(lp.params)
    qkv = gemm(x, lp.Wqkv, bias=lp.bqkv)                                                    # [M, 3H]
    if _fused_attention(L, dh):
        ctxv, lse = K.attn_fwd(qkv, attn_mask, B, heads, L, 1.0 / math.sqrt(dh), p_attn, nctx.seed, sid + 1, Lp=Lp)
        S, Pd = lse, None               # backward recomputes the probabilities from the saved log-sum-exp
    else:
        ctxv, S, Pd = _attention_unfused(qkv, nctx, attn_mask, B, L, Lp, H, heads, dh, sid, p_attn)
    z1 = gemm(ctxv, lp.Wo)
    y1, mean1, rstd1 = K.ln_fwd(z1, lp.g1, lp.be1, bias=lp.bo, residual=x, eps=lp.eps, p_pre=p_hid, seed=nctx.seed,
                                stream_pre=sid + 2)
    u = torch.empty((B * L, lp.W1.shape[0]), dtype=dt, device=dev)
    # u receives gelu'(pre-activation) straight from the epilogue that computes gelu: backward only multiplies
    a = gemm(y1, lp.W1, bias=lp.b1, act=N.ACT_GELU_SAVE_GRAD, preact_out=u)
    z2 = gemm(a, lp.W2)
    y2, mean2, rstd2 = K.ln_fwd(z2, lp.g2, lp.be2, bias=lp.b2, residual=y1, eps=lp.eps, p_pre=p_hid, seed=nctx.seed,
                                stream_pre=sid + 3)
    return y2, (qkv, S, Pd, ctxv, z1, y1, mean1, rstd1, u, a, z2, mean2, rstd2)


def _attention_unfused_bwd(qkv, P, Pd, dctx, nctx, B, L, Lp, H, heads, dh, sid, p_attn):
    """dqkv of the unfused attention: four batched GEMMs around the softmax backward kernel."""
    dt, dev = qkv.dtype, qkv.device
    scale = 1.0 / math.sqrt(dh)
    sS = (L * Lp, heads * L * Lp)
    sQ = (dh, L * 3 * H)
    sC = (dh, L * H)
    dP = torch.empty((B, heads, L, Lp), dtype=dt, device=dev)
    fused = _fused_softmax(L)
    if fused:     # dS = P * (dPe - rowsum(P*dPe)) computed in the epilogue of the dP GEMM
        bgemm(dctx, qkv, dP, M=L, N=L, K=dh, lda=H, ldb=3 * H, ldd=Lp, a_mn=False, b_mn=False, batch0=heads, batch1=B,
              sa=sC, sb=sQ, sd=sS, b_off=2 * H, epi_mode=2, aux=P, p_drop=p_attn, seed=nctx.seed, stream_id=sid + 1)
    else:
        bgemm(dctx, qkv, dP, M=L, N=L, K=dh, lda=H, ldb=3 * H, ldd=Lp, a_mn=False, b_mn=False, batch0=heads, batch1=B,
              sa=sC, sb=sQ, sd=sS, b_off=2 * H)                                               # dP = dctx V^T
    dqkv = torch.empty((B * L, 3 * H), dtype=dt, device=dev)
    bgemm(Pd, dctx, dqkv, M=L, N=dh, K=L, lda=Lp, ldb=H, ldd=3 * H, a_mn=True, b_mn=True, batch0=heads, batch1=B,
          sa=sS, sb=sC, sd=sQ, d_off=2 * H)                                                   # dV = Pd^T dctx
    if not fused:
        K.softmax_bwd_(P, dP, B, heads, L, Lp, p_attn, nctx.seed, sid + 1)                    # dP <- dS
    bgemm(dP, qkv, dqkv, M=L, N=dh, K=L, lda=Lp, ldb=3 * H, ldd=3 * H, a_mn=False, b_mn=True, batch0=heads, batch1=B,
          sa=sS, sb=sQ, sd=sQ, b_off=H, alpha=scale)                                          # dQ = dS K / sqrt(d)
    bgemm(dP, qkv, dqkv, M=L, N=dh, K=L, lda=Lp, ldb=3 * H, ldd=3 * H, a_mn=True, b_mn=True, batch0=heads, batch1=B,
          sa=sS, sb=sQ, sd=sQ, d_off=H, alpha=scale)                                          # dK = dS^T Q / sqrt(d)
    return dqkv


class BertLayerFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, nctx: NativeCtx, lp: LayerPack, attn_mask, B: int, L: int, sid: int):
        H = x.shape[1]
        p_attn = lp.p_attn if nctx.training else 0.0
        p_hid = lp.p_hidden if nctx.training else 0.0
        x = x.contiguous()
        y2, saved = _layer_forward(x, nctx, lp, attn_mask, B, L, sid, p_attn, p_hid)
        ctx.nctx, ctx.lp, ctx.dims, ctx.sid, ctx.ps = nctx, lp, (B, L, _pad8(L), H), sid, (p_attn, p_hid)
        ctx.attn_mask = attn_mask
        ctx.recompute = bool(getattr(nctx, "checkpoint_activations", False))
        if ctx.recompute:
            # activation checkpointing: keep only the layer input; the forward kernels run again in backward and
            # regenerate the very same dropout masks from (seed, call-site id, element index)
            ctx.save_for_backward(x)
        else:
            ctx.save_for_backward(x, *saved)
        return y2

    @staticmethod
    def backward(ctx, dy2):
        nctx, lp, sid = ctx.nctx, ctx.lp, ctx.sid
        B, L, Lp, H = ctx.dims
        p_attn, p_hid = ctx.ps
        if ctx.recompute:
            (x,) = ctx.saved_tensors
            _, saved = _layer_forward(x, nctx, lp, ctx.attn_mask, B, L, sid, p_attn, p_hid)
            qkv, P, Pd, ctxv, z1, y1, mean1, rstd1, u, a, z2, mean2, rstd2 = saved
        else:
            x, qkv, P, Pd, ctxv, z1, y1, mean1, rstd1, u, a, z2, mean2, rstd2 = ctx.saved_tensors
        heads, dh = lp.heads, lp.dh
        ws = nctx.ws
        dt, dev = x.dtype, x.device
        wg = dict(out_dtype=torch.float32, accumulate=True)
        dy2 = dy2.contiguous()

        # ---- FFN -------------------------------------------------------------------------------------
        ds2, dz2 = K.ln_bwd(dy2, z2, lp.g2, lp.be2, mean2, rstd2, _mg(lp.g2), _mg(lp.be2), bias=lp.b2, dbias=_mg(lp.b2),
                            residual=y1, eps=lp.eps, p_pre=p_hid, seed=nctx.seed, stream_pre=sid + 3, want_dz=p_hid > 0, ws=ws)
        if dz2 is None:
            dz2 = ds2
        gemm(dz2, a, a_mn=True, b_mn=True, out=_mg(lp.W2), **wg)                                 # dW2 += dz2^T a
        du = gemm(dz2, lp.W2, b_mn=True, aux=u, aux_mode=N.AUX_MUL)                               # (dz2 W2) * gelu'(pre), u = gelu' 
        nctx.side.colsum_(du, _mg(lp.b1))                                                        # side stream
        gemm(du, y1, a_mn=True, b_mn=True, out=_mg(lp.W1), **wg)                                 # dW1 += du^T y1
        dy1 = gemm(du, lp.W1, b_mn=True, aux=ds2, aux_mode=N.AUX_ADD)                             # + residual branch

        # ---- attention output projection ------------------------------------------------------------------
        ds1, dz1 = K.ln_bwd(dy1, z1, lp.g1, lp.be1, mean1, rstd1, _mg(lp.g1), _mg(lp.be1), bias=lp.bo, dbias=_mg(lp.bo),
                            residual=x, eps=lp.eps, p_pre=p_hid, seed=nctx.seed, stream_pre=sid + 2, want_dz=p_hid > 0, ws=ws)
        if dz1 is None:
            dz1 = ds1
        gemm(dz1, ctxv, a_mn=True, b_mn=True, out=_mg(lp.Wo), **wg)                              # dWo += dz1^T ctx
        dctx = gemm(dz1, lp.Wo, b_mn=True)

        # ---- attention core -------------------------------------------------------------------------------
        scale = 1.0 / math.sqrt(dh)
        if Pd is None:      # fused path: P holds the log-sum-exp saved by attn_fwd
            dqkv = K.attn_bwd(qkv, ctxv, dctx, P, ctx.attn_mask, B, heads, L, scale, p_attn, nctx.seed, sid + 1, Lp=Lp)
        else:
            dqkv = _attention_unfused_bwd(qkv, P, Pd, dctx, nctx, B, L, Lp, H, heads, dh, sid, p_attn)

        # ---- QKV projection -------------------------------------------------------------------------------
        nctx.side.colsum_(dqkv, lp.gbqkv)                                                        # side stream
        gemm(dqkv, x, a_mn=True, b_mn=True, out=lp.gWqkv, **wg)                                  # dWqkv += dqkv^T x
        dx = gemm(dqkv, lp.Wqkv, b_mn=True, aux=ds1, aux_mode=N.AUX_ADD)
        nctx.side.join()                                  # bias gradients are in before the layer is announced
        nctx.sink.ready(lp.params)
        return dx, None, None, None, None, None, None


# ---------------------------------------------------------------------------------------------------
# pooler  (ref: hetseq/bert_modeling.py:506-516): tanh(W h[:,0] + b)
# ---------------------------------------------------------------------------------------------------
class PoolerFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, seq2d, nctx: NativeCtx, pooler, B: int, L: int):
        first = seq2d.view(B, L, -1)[:, 0]                       # [B, H] rows strided by L*H
        nctx.sink.need([pooler.dense_act.weight, pooler.dense_act.bias])
        y = gemm(first, pooler.dense_act.weight, bias=pooler.dense_act.bias, act=N.ACT_TANH)
        ctx.nctx, ctx.pooler, ctx.dims = nctx, pooler, (B, L)
        ctx.save_for_backward(seq2d, y)
        return y

    @staticmethod
    def backward(ctx, dy):
        seq2d, y = ctx.saved_tensors
        B, L = ctx.dims
        W, b = ctx.pooler.dense_act.weight, ctx.pooler.dense_act.bias
        dy = dy.contiguous()
        first = seq2d.view(B, L, -1)[:, 0]
        du = K.act_bwd(dy, y, 2)                                 # dy * (1 - y^2)
        K.colsum_(du, _mg(b), ctx.nctx.ws)
        gemm(du, first, a_mn=True, b_mn=True, out=_mg(W), out_dtype=torch.float32, accumulate=True)
        dfirst = gemm(du, W, b_mn=True)                          # [B, H]
        dseq = torch.zeros_like(seq2d)
        dseq.view(B, L, -1)[:, 0] = dfirst
        ctx.nctx.sink.ready([W, b])
        return dseq, None, None, None, None


# ---------------------------------------------------------------------------------------------------
# pre-training heads + losses  (ref: hetseq/bert_modeling.py:519-581, 894-905)
# ---------------------------------------------------------------------------------------------------
class PretrainLossFn(torch.autograd.Function):
    """loss = CE_mlm(ignore -1, mean over labelled positions) + CE_nsp.

    The MLM head is evaluated only at the labelled positions (gathered on the device, fixed capacity, no host
    sync).  Ignored positions contribute neither to the loss nor to any gradient in the reference either
    (CrossEntropyLoss(ignore_index=-1) averages over non-ignored targets), so the result is identical while the
    [B*L, vocab] logits tensor (500 MB at B=32, L=128) never exists.
    """

    @staticmethod
    def forward(ctx, seq2d, nctx: NativeCtx, model, mlm_labels, nsp_labels, B: int, L: int, cap: int):
        heads = model.cls
        tr = heads.predictions.transform
        emb_w = model.bert.embeddings.word_embeddings.weight
        dec_b = heads.predictions.bias
        V, H = emb_w.shape
        dt, dev = seq2d.dtype, seq2d.device
        labels = mlm_labels.reshape(-1).contiguous()
        nctx.sink.need(list(heads.parameters()) + list(model.bert.pooler.parameters()))

        idx, lab, cnt = K.compact_rows(labels, cap, ignore=-1)
        hsel = K.gather_rows(seq2d, idx)                                                   # [cap, H]
        u = torch.empty((cap, H), dtype=dt, device=dev)
        t = gemm(hsel, tr.dense_act.weight, bias=tr.dense_act.bias, act=N.ACT_GELU, preact_out=u)
        tn, mean, rstd = K.ln_fwd(t, tr.LayerNorm.weight, tr.LayerNorm.bias, eps=tr.LayerNorm.variance_epsilon)
        pad = 16 // seq2d.element_size()
        ldv = (V + pad - 1) // pad * pad
        logits = torch.empty((cap, ldv), dtype=dt, device=dev)[:, :V]
        gemm(tn, emb_w, bias=dec_b, out=logits)
        lse, row_loss = K.ce_fwd(logits, lab, ignore=-1)
        mlm_loss, inv_cnt, _ = K.ce_finalize(row_loss, lab, ignore=-1)

        # pooler + NSP
        first = seq2d.view(B, L, -1)[:, 0]
        pw, pb = model.bert.pooler.dense_act.weight, model.bert.pooler.dense_act.bias
        pooled = gemm(first, pw, bias=pb, act=N.ACT_TANH)
        sw, sb = heads.seq_relationship.weight, heads.seq_relationship.bias
        nsp_logits = K.head_fwd(pooled, sw, sb)
        nsp_lab = nsp_labels.reshape(-1).contiguous()
        nsp_rows, nsp_eff = K.head_ce(nsp_logits, nsp_lab, None, ignore=-1)
        nsp_loss, nsp_inv, _ = K.ce_finalize(nsp_rows, nsp_eff, ignore=-1)

        loss = (mlm_loss + nsp_loss).to(torch.float32).reshape(())
        ctx.nctx, ctx.model, ctx.dims = nctx, model, (B, L, cap, V, H)
        ctx.save_for_backward(seq2d, idx, lab, hsel, u, t, tn, mean, rstd, logits, lse, inv_cnt, pooled, nsp_logits,
                              nsp_eff, nsp_inv)
        ctx.mark_non_differentiable(cnt)
        return loss, cnt

    @staticmethod
    def backward(ctx, gloss, _gcnt):
        (seq2d, idx, lab, hsel, u, t, tn, mean, rstd, logits, lse, inv_cnt, pooled, nsp_logits, nsp_eff,
         nsp_inv) = ctx.saved_tensors
        nctx, model = ctx.nctx, ctx.model
        B, L, cap, V, H = ctx.dims
        ws = nctx.ws
        heads = model.cls
        tr = heads.predictions.transform
        emb_w = model.bert.embeddings.word_embeddings.weight
        dec_b = heads.predictions.bias
        dt = seq2d.dtype
        wg = dict(out_dtype=torch.float32, accumulate=True)
        g = gloss.reshape(1).to(torch.float32).contiguous()

        # ---- MLM branch ------------------------------------------------------------------------------
        dlog = K.ce_bwd_(logits, lab, lse, gout=g, inv_count=inv_cnt, ignore=-1)            # in place
        if V % 4 == 0:
            K.colsum_(dlog, _mg(dec_b), ws)
        else:
            _mg(dec_b).add_(dlog.float().sum(0))                                            # odd vocab: tiny torch reduction
        gemm(dlog, tn, a_mn=True, b_mn=True, out=_mg(emb_w), **wg)                          # tied decoder: dE += dlog^T tn
        dtn = gemm(dlog, emb_w, b_mn=True)                                                  # [cap, H]
        dt_, _ = K.ln_bwd(dtn, t, tr.LayerNorm.weight, tr.LayerNorm.bias, mean, rstd, _mg(tr.LayerNorm.weight),
                          _mg(tr.LayerNorm.bias), eps=tr.LayerNorm.variance_epsilon, ws=ws)
        du = K.act_bwd(dt_, u, 1)                                                           # dt * gelu'(u)
        K.colsum_(du, _mg(tr.dense_act.bias), ws)
        gemm(du, hsel, a_mn=True, b_mn=True, out=_mg(tr.dense_act.weight), **wg)
        dhsel = gemm(du, tr.dense_act.weight, b_mn=True)                                    # [cap, H]
        dseq = torch.zeros_like(seq2d)
        K.scatter_rows(dhsel, idx, dseq, accumulate=False)

        # ---- NSP branch ------------------------------------------------------------------------------
        sw, sb = heads.seq_relationship.weight, heads.seq_relationship.bias
        pw, pb = model.bert.pooler.dense_act.weight, model.bert.pooler.dense_act.bias
        dnl = K.head_ce_bwd(nsp_logits, nsp_eff, gout=g, inv_count=nsp_inv, ignore=-1)       # [B, 2] fp32
        dpooled = torch.empty_like(pooled)
        K.head_bwd(dnl, pooled, sw, dx=dpooled, accumulate_dx=False, dw=_mg(sw), db=_mg(sb))
        dup = K.act_bwd(dpooled, pooled, 2)                                                  # [B, H] tanh'
        K.colsum_(dup, _mg(pb), ws)
        first = seq2d.view(B, L, -1)[:, 0]
        gemm(dup, first, a_mn=True, b_mn=True, out=_mg(pw), **wg)
        dfirst = gemm(dup, pw, b_mn=True)
        dseq.view(B, L, -1)[:, 0] += dfirst
        # the decoder weight is the word-embedding table (tied): its gradient is final only after the embedding
        # backward has scattered the token-lookup part into it, so EmbeddingsFn.backward announces it, not this block
        nctx.sink.ready([p for p in heads.parameters() if p is not emb_w] + [pw, pb])
        return dseq, None, None, None, None, None, None, None


# ---------------------------------------------------------------------------------------------------
# token-classification head + loss  (ref: hetseq/bert_modeling.py:1221-1245)
# ---------------------------------------------------------------------------------------------------
class TokenClsLossFn(torch.autograd.Function):
    """dropout -> Linear(H, num_labels) -> CE over positions with attention_mask == 1 (ignore_index -100),
    with the active-position selection done by label rewriting on the device (no boolean-mask gather, no
    host sync)."""

    @staticmethod
    def forward(ctx, seq2d, nctx: NativeCtx, model, labels, attention_mask, sid: int):
        p = model.dropout.p if nctx.training else 0.0
        w, b = model.classifier.weight, model.classifier.bias
        nctx.sink.need([w, b])
        x = K.dropout(seq2d.contiguous(), p, nctx.seed, sid) if p > 0 else seq2d.contiguous()
        logits = K.head_fwd(x, w, b)                                                         # fp32 [M, C]
        lab = labels.reshape(-1).contiguous()
        act = attention_mask.reshape(-1).contiguous() if attention_mask is not None else None
        rows, eff = K.head_ce(logits, lab, act, ignore=-100)
        loss, inv, _ = K.ce_finalize(rows, eff, ignore=-100)
        ctx.nctx, ctx.model, ctx.sid, ctx.p = nctx, model, sid, p
        ctx.save_for_backward(x, logits, eff, inv)
        ctx.mark_non_differentiable(logits)
        return loss.reshape(()), logits

    @staticmethod
    def backward(ctx, gloss, _glogits):
        x, logits, eff, inv = ctx.saved_tensors
        model, nctx = ctx.model, ctx.nctx
        w, b = model.classifier.weight, model.classifier.bias
        g = gloss.reshape(1).to(torch.float32).contiguous()
        dl = K.head_ce_bwd(logits, eff, gout=g, inv_count=inv, ignore=-100)
        dx = torch.empty_like(x)
        K.head_bwd(dl, x, w, dx=dx, accumulate_dx=False, dw=_mg(w), db=_mg(b))
        if ctx.p > 0:
            dx = K.dropout(dx, ctx.p, nctx.seed, ctx.sid)
        nctx.sink.ready([w, b])
        return dx, None, None, None, None, None


# ---------------------------------------------------------------------------------------------------
# helpers used by bert_modeling
# ---------------------------------------------------------------------------------------------------
def native_supported(config, L: int) -> Optional[str]:
    """None if the kernels cover this configuration, else the reason they do not."""
    H, heads = config.hidden_size, config.num_attention_heads
    if H % 64 or H > 1024:
        return "hidden_size must be a multiple of 64 and <= 1024"
    if config.intermediate_size % 64:
        return "intermediate_size must be a multiple of 64"
    if (H // heads) % 8 or (H // heads) > 256:
        return "head size must be a multiple of 8"
    if L > 512:
        return "sequence length > 512"
    if getattr(config, "hidden_act", "gelu") != "gelu":
        return "only gelu is fused"
    return None


def encoder_forward(bert, nctx: NativeCtx, input_ids, token_type_ids, attention_mask) -> torch.Tensor:
    """Embeddings + all encoder layers; returns the final hidden states as [B*L, H]."""
    B, L = input_ids.shape
    packs: List[LayerPack] = bert._native_packs
    anchor = bert.embeddings.word_embeddings.weight
    mask = attention_mask.contiguous() if attention_mask is not None else None
    if mask is not None and mask.dtype != torch.int64:
        mask = mask.to(torch.int64)
    h = EmbeddingsFn.apply(anchor, nctx, bert.embeddings, input_ids, token_type_ids, L, 1 << 20)
    for i, lp in enumerate(packs):
        h = BertLayerFn.apply(h, nctx, lp, mask, B, L, (i + 1) * 8)
    return h
