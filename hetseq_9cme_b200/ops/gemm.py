"""Python front-end of the tcgen05 GEMM (``csrc/gemm_sm100.cu``).

``gemm(a, b, ...)`` computes ``D[M,N] = epilogue(alpha * A @ B^T-like)`` where each operand may be
stored K-major or MN-major; see ``include/b200seq.h``.  Replaces every cuBLAS call behind ``nn.Linear``
in the reference (``hetseq/bert_modeling.py:166-172,352-354,388,424,515,526,544``).
"""
from __future__ import annotations

import ctypes
from typing import Optional

import torch

from . import native

_DT = {torch.float32: 4, torch.bfloat16: 2}

# ---- strict-fp32 policy ------------------------------------------------------------------------------------------
# The reference computes in fp32 (hetseq/options.py:11).  `--precision tf32` feeds fp32 storage to the tensor core
# as TF32 (10-bit mantissa); `--precision fp32` keeps the tensor core but splits every fp32 operand into two
# TF32-exact halves and issues three products (hi.hi + hi.lo + lo.hi, fp32 accumulation): about 2^-22 relative error
# per product instead of 2^-11, at ~3.5x the TF32 cost.  Epilogues run after the three partial GEMMs.
_POLICY = {"strict_fp32": False}


def set_strict_fp32(on: bool) -> None:
    _POLICY["strict_fp32"] = bool(on)


def strict_fp32() -> bool:
    return _POLICY["strict_fp32"]


def _split(t: torch.Tensor):
    from . import kernels
    return kernels.split_tf32(t if t.is_contiguous() else t.contiguous())


def _gelu_grad(x: torch.Tensor) -> torch.Tensor:
    cdf = 0.5 * (1.0 + torch.erf(x * 0.7071067811865476))
    return cdf + x * torch.exp(-0.5 * x * x) * 0.3989422804014327


def _gemm_strict(a, b, *, a_mn, b_mn, out, out_dtype, bias, act, aux, aux_mode, preact_out, accumulate, alpha, bn, splits):
    ra, ca = a.shape
    rb, cb = b.shape
    M = ca if a_mn else ra
    N = cb if b_mn else rb
    a_hi, a_lo = _split(a)
    b_hi, b_lo = _split(b)
    plain = bias is None and act == native.ACT_NONE and aux is None and preact_out is None
    kw = dict(a_mn=a_mn, b_mn=b_mn, alpha=alpha, bn=bn, splits=splits)
    if plain:
        if out is None:
            out = torch.empty((M, N), dtype=torch.float32, device=a.device)
        _gemm_raw(a_hi, b_lo, out=out, accumulate=accumulate, **kw)
        _gemm_raw(a_lo, b_hi, out=out, accumulate=True, **kw)
        _gemm_raw(a_hi, b_hi, out=out, accumulate=True, **kw)
        return out
    if accumulate:
        raise ValueError("gemm(strict fp32): accumulate with an epilogue is not supported")
    acc = torch.empty((M, N), dtype=torch.float32, device=a.device)
    _gemm_raw(a_hi, b_lo, out=acc, accumulate=False, **kw)
    _gemm_raw(a_lo, b_hi, out=acc, accumulate=True, **kw)
    _gemm_raw(a_hi, b_hi, out=acc, accumulate=True, **kw)
    v = acc if bias is None else acc.add_(bias)
    if act == native.ACT_GELU_SAVE_GRAD:
        preact_out.copy_(_gelu_grad(v))
        v = torch.nn.functional.gelu(v)
    else:
        if preact_out is not None:
            preact_out.copy_(v)
        if act == native.ACT_GELU:
            v = torch.nn.functional.gelu(v)
        elif act == native.ACT_TANH:
            v = torch.tanh(v)
    if aux_mode == native.AUX_ADD:
        v = v + aux
    elif aux_mode == native.AUX_MUL:
        v = v * aux
    elif aux_mode == native.AUX_MUL_DGELU:
        v = v * _gelu_grad(aux)
    elif aux_mode == native.AUX_MUL_DTANH:
        v = v * (1.0 - aux * aux)
    if out is None:
        return v
    out.copy_(v)
    return out


def _stream_ptr() -> int:
    return torch.cuda.current_stream().cuda_stream


def _rowmajor_2d(t: torch.Tensor, name: str):
    if t.dim() != 2 or t.stride(1) != 1:
        raise ValueError(f"{name}: expected a 2-D tensor with unit inner stride, got shape={tuple(t.shape)} stride={t.stride()}")
    return t.shape[0], t.shape[1], t.stride(0)


def gemm(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
         out: Optional[torch.Tensor] = None, out_dtype: Optional[torch.dtype] = None,
         bias: Optional[torch.Tensor] = None, act: int = native.ACT_NONE,
         aux: Optional[torch.Tensor] = None, aux_mode: int = native.AUX_NONE,
         preact_out: Optional[torch.Tensor] = None, accumulate: bool = False, alpha: float = 1.0,
         bn: int = 0, splits: int = 0) -> torch.Tensor:
    """``_gemm_raw`` under the precision policy: fp32 operands go through the 3xTF32 split when ``--precision fp32``."""
    if _POLICY["strict_fp32"] and a.dtype == torch.float32 and b.dtype == torch.float32:
        return _gemm_strict(a, b, a_mn=a_mn, b_mn=b_mn, out=out, out_dtype=out_dtype, bias=bias, act=act, aux=aux,
                            aux_mode=aux_mode, preact_out=preact_out, accumulate=accumulate, alpha=alpha, bn=bn, splits=splits)
    return _gemm_raw(a, b, a_mn=a_mn, b_mn=b_mn, out=out, out_dtype=out_dtype, bias=bias, act=act, aux=aux, aux_mode=aux_mode,
                     preact_out=preact_out, accumulate=accumulate, alpha=alpha, bn=bn, splits=splits)


def _gemm_raw(a: torch.Tensor, b: torch.Tensor, *, a_mn: bool = False, b_mn: bool = False,
              out: Optional[torch.Tensor] = None, out_dtype: Optional[torch.dtype] = None,
              bias: Optional[torch.Tensor] = None, act: int = native.ACT_NONE,
              aux: Optional[torch.Tensor] = None, aux_mode: int = native.AUX_NONE,
              preact_out: Optional[torch.Tensor] = None, accumulate: bool = False, alpha: float = 1.0,
              bn: int = 0, splits: int = 0) -> torch.Tensor:
    """D = epilogue(alpha * sum_k A(m,k) B(n,k)).

    a: [M,K] (a_mn=False) or [K,M] (a_mn=True);  b: [N,K] (b_mn=False) or [K,N] (b_mn=True).
    fp32 operands run on kind::tf32, bf16 operands on kind::f16; output dtype defaults to the operand
    dtype (bf16 operands may request fp32 output, used for weight gradients).
    ``preact_out`` receives the value before the activation — or, with ``act=ACT_GELU_SAVE_GRAD``, gelu'(value), which
    the backward GEMM then applies with ``aux_mode=AUX_MUL`` instead of evaluating erf / exp again.
    """
    if a.dtype != b.dtype or a.dtype not in _DT:
        raise ValueError(f"gemm: unsupported operand dtypes {a.dtype}/{b.dtype}")
    ra, ca, lda = _rowmajor_2d(a, "a")
    rb, cb, ldb = _rowmajor_2d(b, "b")
    M, K = (ca, ra) if a_mn else (ra, ca)
    N, Kb = (cb, rb) if b_mn else (rb, cb)
    if K != Kb:
        raise ValueError(f"gemm: reduction mismatch {K} vs {Kb}")
    odt = out_dtype or (out.dtype if out is not None else a.dtype)
    if a.dtype == torch.float32:
        if odt != torch.float32:
            raise ValueError("gemm: fp32 operands produce fp32 output")
        dt = native.F32_TF32
    else:
        dt = native.BF16 if odt == torch.bfloat16 else native.BF16_F32OUT
    if out is None:
        if accumulate:
            raise ValueError("gemm: accumulate=True needs an explicit out tensor")
        pad = 16 // _DT[odt]
        ldd = (N + pad - 1) // pad * pad
        out = torch.empty((M, ldd), dtype=odt, device=a.device)[:, :N] if ldd != N else torch.empty((M, N), dtype=odt, device=a.device)
    ro, co, ldd = _rowmajor_2d(out, "out")
    if (ro, co) != (M, N) or out.dtype != odt:
        raise ValueError(f"gemm: out has shape {tuple(out.shape)}/{out.dtype}, expected {(M, N)}/{odt}")
    d = native.GemmDesc()
    d.dtype, d.M, d.N, d.K = dt, M, N, K
    d.A, d.lda, d.a_mn = a.data_ptr(), lda, int(a_mn)
    d.B, d.ldb, d.b_mn = b.data_ptr(), ldb, int(b_mn)
    d.D, d.ldd = out.data_ptr(), ldd
    if preact_out is not None:
        rp, cp, ldp = _rowmajor_2d(preact_out, "preact_out")
        if (rp, cp, ldp) != (M, N, ldd) or preact_out.dtype != odt:
            raise ValueError("gemm: preact_out must match out in shape, pitch and dtype")
        d.D2 = preact_out.data_ptr()
    else:
        d.D2 = None
    if bias is not None:
        if bias.dtype != odt or bias.numel() != N or not bias.is_contiguous():
            raise ValueError("gemm: bias must be a contiguous [N] tensor of the output dtype")
        d.bias = bias.data_ptr()
    else:
        d.bias = None
    if aux is not None:
        rx, cx, ldx = _rowmajor_2d(aux, "aux")
        if (rx, cx) != (M, N) or aux.dtype != odt:
            raise ValueError("gemm: aux must be [M,N] of the output dtype")
        d.aux, d.ldaux = aux.data_ptr(), ldx
    else:
        d.aux, d.ldaux = None, 0
        if aux_mode != native.AUX_NONE:
            raise ValueError("gemm: aux_mode without aux")
    d.act, d.aux_mode, d.accumulate, d.alpha = act, aux_mode, int(accumulate), float(alpha)
    d.bn, d.splits = bn, splits
    d.batch0 = d.batch1 = 0
    d.epi_mode = 0
    native.check(native.cuda_lib().b200_gemm(ctypes.byref(d), _stream_ptr()), "b200_gemm")
    return out


def bgemm(a_ptr_t: torch.Tensor, b_ptr_t: torch.Tensor, out_t: torch.Tensor, *, M: int, N: int, K: int,
          lda: int, ldb: int, ldd: int, a_mn: bool, b_mn: bool, batch0: int, batch1: int,
          sa=(0, 0), sb=(0, 0), sd=(0, 0), a_off: int = 0, b_off: int = 0, d_off: int = 0, alpha: float = 1.0,
          out_fp32: bool = False, bn: int = 0, epi_mode: int = 0, attn_mask=None, p_drop: float = 0.0, seed=None,
          stream_id: int = 0, d2: Optional[torch.Tensor] = None, aux: Optional[torch.Tensor] = None,
          accumulate: bool = False) -> None:
    """Batched GEMM over (batch1 x batch0) problems addressed by element strides — used for attention, where
    batch0 = head and batch1 = sequence.  ``*_off`` are element offsets into the given storage tensors."""
    if _POLICY["strict_fp32"] and a_ptr_t.dtype == torch.float32 and epi_mode == 0 and not accumulate:
        # strict fp32: the storage tensors are split wholesale (offsets and strides stay valid), three batched products
        (a_hi, a_lo), (b_hi, b_lo) = _split(a_ptr_t), _split(b_ptr_t)
        kw = dict(M=M, N=N, K=K, lda=lda, ldb=ldb, ldd=ldd, a_mn=a_mn, b_mn=b_mn, batch0=batch0, batch1=batch1, sa=sa, sb=sb,
                  sd=sd, a_off=a_off, b_off=b_off, d_off=d_off, alpha=alpha, out_fp32=out_fp32, bn=bn)
        _POLICY["strict_fp32"] = False
        try:
            bgemm(a_hi, b_lo, out_t, accumulate=False, **kw)
            bgemm(a_lo, b_hi, out_t, accumulate=True, **kw)
            bgemm(a_hi, b_hi, out_t, accumulate=True, **kw)
        finally:
            _POLICY["strict_fp32"] = True
        return
    es = a_ptr_t.element_size()
    d = native.GemmDesc()
    if a_ptr_t.dtype == torch.float32:
        d.dtype = native.F32_TF32
    else:
        d.dtype = native.BF16_F32OUT if out_fp32 else native.BF16
    d.M, d.N, d.K = M, N, K
    d.A, d.lda, d.a_mn = a_ptr_t.data_ptr() + a_off * es, lda, int(a_mn)
    d.B, d.ldb, d.b_mn = b_ptr_t.data_ptr() + b_off * es, ldb, int(b_mn)
    d.D, d.ldd = out_t.data_ptr() + d_off * out_t.element_size(), ldd
    d.D2 = d2.data_ptr() if d2 is not None else None
    d.bias = None
    d.aux = aux.data_ptr() if aux is not None else None
    d.ldaux = ldd if aux is not None else 0
    d.act = d.aux_mode = 0
    d.accumulate = int(accumulate)
    d.epi_mode = epi_mode
    d.attn_mask = attn_mask.data_ptr() if attn_mask is not None else None
    d.p_drop = float(p_drop)
    d.seed_dev = seed.data_ptr() if seed is not None else None
    d.stream_id = int(stream_id)
    d.alpha = float(alpha)
    d.bn, d.splits = bn, 0
    d.batch0, d.batch1 = batch0, batch1
    d.sa0, d.sa1 = sa
    d.sb0, d.sb1 = sb
    d.sd0, d.sd1 = sd
    native.check(native.cuda_lib().b200_gemm(ctypes.byref(d), _stream_ptr()), "b200_gemm(batched)")
