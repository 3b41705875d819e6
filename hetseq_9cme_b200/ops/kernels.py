"""Thin Python wrappers over the C ABI in ``include/b200seq.h`` (everything except the GEMM, see gemm.py).

Each wrapper validates shapes/dtypes, passes raw pointers plus the current CUDA stream, and raises
``NativeError`` on failure.  No wrapper allocates except for explicitly requested outputs; scratch space
comes from :class:`Workspace` (one per device, reused by every call on the training stream).
"""
from __future__ import annotations

import os

import ctypes
from ctypes import POINTER, byref, c_double, c_float, c_int, c_longlong, c_ulonglong, c_void_p
from typing import Optional

import torch

from . import native
from .native import check, cuda_lib

FP32, BF16 = 0, 1
_LN_TWO_PASS = bool(__import__('os').environ.get('B200_LN_TWO_PASS'))


def dtype_code(t: torch.dtype) -> int:
    if t == torch.float32:
        return FP32
    if t == torch.bfloat16:
        return BF16
    raise ValueError(f"unsupported tensor dtype {t}")


def _ptr(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream():
    return torch.cuda.current_stream().cuda_stream


class AdamArgs(ctypes.Structure):
    _fields_ = [("lr", c_float), ("beta1", c_float), ("beta2", c_float), ("eps", c_float), ("weight_decay", c_float),
                ("step", c_int), ("grad_scale_mult", c_float), ("max_norm", c_float),
                ("norm_sq", c_void_p), ("sample_size", c_void_p), ("stats_out", c_void_p), ("zero_grad", c_int),
                ("sample_size_f32", c_void_p), ("hyper_dev", c_void_p)]


class AdadeltaArgs(ctypes.Structure):
    _fields_ = [("lr", c_float), ("rho", c_float), ("eps", c_float), ("weight_decay", c_float),
                ("grad_scale_mult", c_float), ("max_norm", c_float),
                ("norm_sq", c_void_p), ("sample_size", c_void_p), ("stats_out", c_void_p), ("zero_grad", c_int),
                ("sample_size_f32", c_void_p), ("hyper_dev", c_void_p)]


class LnArgs(ctypes.Structure):
    _fields_ = [("dtype", c_int), ("rows", c_int), ("cols", c_int),
                ("x", c_void_p), ("ldx", c_longlong),
                ("bias", c_void_p),
                ("residual", c_void_p), ("ldr", c_longlong),
                ("gamma", c_void_p), ("beta", c_void_p),
                ("eps", c_float), ("p_pre", c_float), ("p_post", c_float),
                ("seed_dev", c_void_p), ("stream_pre", c_ulonglong), ("stream_post", c_ulonglong),
                ("y", c_void_p), ("ldy", c_longlong),
                ("mean", c_void_p), ("rstd", c_void_p),
                ("dy", c_void_p), ("lddy", c_longlong),
                ("ds", c_void_p), ("ldds", c_longlong),
                ("dz", c_void_p), ("lddz", c_longlong),
                ("dgamma", c_void_p), ("dbeta", c_void_p), ("dbias", c_void_p),
                ("scratch", c_void_p), ("scratch_bytes", c_longlong)]


class AttnArgs(ctypes.Structure):
    _fields_ = [("dtype", c_int), ("B", c_int), ("heads", c_int), ("L", c_int), ("Lp", c_int),
                ("scale", c_float), ("p_drop", c_float),
                ("qkv", c_void_p), ("ld_qkv", c_longlong),
                ("attn_mask", c_void_p), ("seed_dev", c_void_p), ("stream_id", c_ulonglong),
                ("out", c_void_p), ("ld_out", c_longlong),
                ("lse", c_void_p), ("delta", c_void_p), ("o", c_void_p), ("dout", c_void_p), ("ld_o", c_longlong)]


_declared = False


def _lib():
    global _declared
    lib = cuda_lib()
    if not _declared:
        I, LL, P, F, ULL = c_int, c_longlong, c_void_p, c_float, c_ulonglong
        sig = {
            "b200_sumsq_scratch_bytes": ([], LL),
            "b200_sumsq_f32": ([P, LL, P, P, LL, I, P], I),
            "b200_adam_flat": ([P, P, P, P, P, LL, POINTER(AdamArgs), P], I),
            "b200_adadelta_flat": ([P, P, P, P, P, LL, POINTER(AdadeltaArgs), P], I),
            "b200_scale_f32": ([P, LL, F, P, P], I),
            "b200_cast_f32_to_bf16": ([P, P, LL, P], I),
            "b200_split_tf32": ([P, P, P, LL, P], I),
            "b200_ln_bwd_scratch_bytes": ([I], LL),
            "b200_ln_fwd": ([POINTER(LnArgs), P], I),
            "b200_ln_bwd": ([POINTER(LnArgs), P], I),
            "b200_embed_sum": ([I, P, P, P, P, P, P, I, I, I, I, I, P], I),
            "b200_embed_scatter": ([I, P, P, P, P, P, P, I, I, I, I, I, P], I),
            "b200_compact_rows": ([P, P, I, I, I, P, P, P, P], I),
            "b200_gather_rows": ([I, P, LL, P, I, I, P, P], I),
            "b200_scatter_rows": ([I, P, P, I, I, P, LL, I, P], I),
            "b200_ce_fwd": ([I, P, LL, P, I, I, I, P, P, P], I),
            "b200_ce_bwd": ([I, P, LL, P, I, I, I, P, P, P, P], I),
            "b200_ce_finalize": ([P, P, I, I, P, P, P, P], I),
            "b200_head_fwd": ([I, P, LL, P, P, I, I, I, P, P], I),
            "b200_head_ce": ([P, P, P, I, I, I, P, P, P], I),
            "b200_head_ce_bwd": ([P, P, I, I, I, P, P, P, P], I),
            "b200_head_bwd": ([I, P, P, LL, P, I, I, I, P, LL, I, P, P, P], I),
            "b200_colsum_scratch_bytes": ([I, I], LL),
            "b200_colsum": ([I, P, LL, I, I, P, P, LL, P], I),
            "b200_softmax_fwd": ([I, P, P, P, I, I, I, I, F, P, ULL, P], I),
            "b200_softmax_bwd": ([I, P, P, I, I, I, I, F, P, ULL, P], I),
            "b200_assemble_bert_batch": ([P, I, I, I, P, P, P, P, P, P], I),
            "b200_attn_fwd": ([POINTER(AttnArgs), P], I),
            "b200_attn_bwd": ([POINTER(AttnArgs), P], I),
            "b200_dropout": ([I, P, P, LL, F, P, ULL, P], I),
            "b200_act_bwd": ([I, P, P, P, LL, I, P], I),
        }
        for name, (args, res) in sig.items():
            fn = getattr(lib, name)
            fn.argtypes = args
            fn.restype = res
        _declared = True
    return lib


# ---------------------------------------------------------------------------------------------------
# workspace
# ---------------------------------------------------------------------------------------------------
class Workspace:
    """Per-device scratch owned by the framework: reduction partials, the dropout seed word, and the small
    device/pinned mailboxes the fused optimizer step uses.  All users run on one stream, in order."""

    _per_device = {}

    def __init__(self, device: torch.device):
        lib = _lib()
        self.device = device
        self.sumsq_scratch = torch.zeros(int(lib.b200_sumsq_scratch_bytes()) // 4 + 4, dtype=torch.float32, device=device)
        self._ln_scratch = None
        self._colsum_scratch = None
        self.seed = torch.zeros(1, dtype=torch.int64, device=device)       # dropout seed word (device resident)
        self._seed_host = torch.zeros(1, dtype=torch.int64).pin_memory()
        self.norm_sq = torch.zeros(1, dtype=torch.float32, device=device)
        self.step_stats = torch.zeros(4, dtype=torch.float32, device=device)  # gnorm, scale, clipped

    @classmethod
    def get(cls, device=None) -> "Workspace":
        device = torch.device(device if device is not None else torch.cuda.current_device())
        if device.type != "cuda":
            raise ValueError("Workspace is CUDA only")
        if device.index is None:
            device = torch.device("cuda", torch.cuda.current_device())
        ws = cls._per_device.get(device.index)
        if ws is None:
            ws = cls._per_device[device.index] = Workspace(device)
        return ws

    def ln_scratch(self, cols: int) -> torch.Tensor:
        need = int(_lib().b200_ln_bwd_scratch_bytes(cols)) // 4
        if self._ln_scratch is None or self._ln_scratch.numel() < need:
            self._ln_scratch = torch.empty(need, dtype=torch.float32, device=self.device)
        return self._ln_scratch

    def colsum_scratch(self, rows: int, cols: int) -> torch.Tensor:
        need = int(_lib().b200_colsum_scratch_bytes(rows, cols)) // 4
        if self._colsum_scratch is None or self._colsum_scratch.numel() < need:
            self._colsum_scratch = torch.empty(max(need, 64 * 4096), dtype=torch.float32, device=self.device)
        return self._colsum_scratch

    def set_seed(self, seed: int):
        """Stage this step's dropout seed on the device (async pinned H2D; graph-capturable)."""
        self._seed_host[0] = int(seed) & 0x7FFFFFFFFFFFFFFF
        self.seed.copy_(self._seed_host, non_blocking=True)


# ---------------------------------------------------------------------------------------------------
# optimizer pieces
# ---------------------------------------------------------------------------------------------------
def sumsq(g: torch.Tensor, out: torch.Tensor, accumulate: bool = False, ws: Optional[Workspace] = None):
    ws = ws or Workspace.get(g.device)
    assert g.dtype == torch.float32 and g.is_contiguous() and out.dtype == torch.float32
    sc = ws.sumsq_scratch
    check(_lib().b200_sumsq_f32(g.data_ptr(), g.numel(), out.data_ptr(), sc.data_ptr(), sc.numel() * 4, int(accumulate),
                                _stream()), "b200_sumsq_f32")
    return out


def make_adam_args(*, lr, beta1, beta2, eps, weight_decay, step, grad_scale_mult=1.0, max_norm=0.0, norm_sq=None,
                   sample_size=None, stats_out=None, zero_grad=True, hyper_dev=None) -> AdamArgs:
    ss64 = sample_size if (sample_size is not None and sample_size.dtype == torch.float64) else None
    ss32 = sample_size if (sample_size is not None and sample_size.dtype == torch.float32) else None
    return AdamArgs(lr, beta1, beta2, eps, weight_decay, int(step), grad_scale_mult, max_norm, _ptr(norm_sq), _ptr(ss64),
                    _ptr(stats_out), int(zero_grad), _ptr(ss32), _ptr(hyper_dev))


def adam_flat(p, g, m, v, n, *, p_bf16=None, **kw):
    a = make_adam_args(**kw)
    check(_lib().b200_adam_flat(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), _ptr(p_bf16), int(n), byref(a),
                                _stream()), "b200_adam_flat")


def adadelta_flat(p, g, sq, acc, n, *, lr, rho, eps, weight_decay, grad_scale_mult=1.0, max_norm=0.0, norm_sq=None,
                  sample_size=None, stats_out=None, zero_grad=True, p_bf16=None, hyper_dev=None):
    ss64 = sample_size if (sample_size is not None and sample_size.dtype == torch.float64) else None
    ss32 = sample_size if (sample_size is not None and sample_size.dtype == torch.float32) else None
    a = AdadeltaArgs(lr, rho, eps, weight_decay, grad_scale_mult, max_norm, _ptr(norm_sq), _ptr(ss64),
                     _ptr(stats_out), int(zero_grad), _ptr(ss32), _ptr(hyper_dev))
    check(_lib().b200_adadelta_flat(p.data_ptr(), g.data_ptr(), sq.data_ptr(), acc.data_ptr(), _ptr(p_bf16), int(n),
                                    byref(a), _stream()), "b200_adadelta_flat")


def scale_(x: torch.Tensor, c: float, c_dev: Optional[torch.Tensor] = None):
    assert x.dtype == torch.float32 and x.is_contiguous()
    check(_lib().b200_scale_f32(x.data_ptr(), x.numel(), float(c), _ptr(c_dev), _stream()), "b200_scale_f32")
    return x


def split_tf32(x: torch.Tensor):
    """(hi, lo) TF32-exact halves of a contiguous fp32 tensor (strict-fp32 GEMM policy, see ops/gemm.py)."""
    assert x.dtype == torch.float32 and x.is_contiguous()
    hi, lo = torch.empty_like(x), torch.empty_like(x)
    check(_lib().b200_split_tf32(x.data_ptr(), hi.data_ptr(), lo.data_ptr(), x.numel(), _stream()), "b200_split_tf32")
    return hi, lo


def cast_to_bf16(src: torch.Tensor, dst: torch.Tensor):
    assert src.dtype == torch.float32 and dst.dtype == torch.bfloat16 and src.numel() == dst.numel()
    check(_lib().b200_cast_f32_to_bf16(src.data_ptr(), dst.data_ptr(), src.numel(), _stream()), "b200_cast_f32_to_bf16")
    return dst


# ---------------------------------------------------------------------------------------------------
# layer norm
# ---------------------------------------------------------------------------------------------------
def _ln_base(x, bias, residual, gamma, beta, eps, p_pre, p_post, seed, stream_pre, stream_post) -> LnArgs:
    rows, cols = x.shape
    a = LnArgs()
    a.dtype, a.rows, a.cols = dtype_code(x.dtype), rows, cols
    a.x, a.ldx = x.data_ptr(), x.stride(0)
    a.bias = _ptr(bias)
    a.residual, a.ldr = (_ptr(residual), residual.stride(0)) if residual is not None else (None, 0)
    a.gamma, a.beta = gamma.data_ptr(), beta.data_ptr()
    a.eps, a.p_pre, a.p_post = eps, p_pre, p_post
    a.seed_dev = _ptr(seed)
    a.stream_pre, a.stream_post = stream_pre, stream_post
    return a


def ln_fwd(x, gamma, beta, *, bias=None, residual=None, eps=1e-12, p_pre=0.0, p_post=0.0, seed=None, stream_pre=0,
           stream_post=0):
    """y = drop_post(LN(drop_pre(x + bias) + residual)); returns (y, mean, rstd)."""
    rows, cols = x.shape
    y = torch.empty((rows, cols), dtype=x.dtype, device=x.device)
    mean = torch.empty(rows, dtype=torch.float32, device=x.device)
    rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
    a = _ln_base(x, bias, residual, gamma, beta, eps, p_pre, p_post, seed, stream_pre, stream_post)
    a.y, a.ldy = y.data_ptr(), y.stride(0)
    a.mean, a.rstd = mean.data_ptr(), rstd.data_ptr()
    check(_lib().b200_ln_fwd(byref(a), _stream()), "b200_ln_fwd")
    return y, mean, rstd


def ln_config(pipe: int = -2, fwd_ctas: int = 0, bwd_ctas: int = 0) -> int:
    """Select the streaming (cp.async) LayerNorm kernels: ``pipe`` bit 0 = forward, bit 1 = backward, -1 = automatic
    (streaming backward for fp32 storage); ``*_ctas`` = CTAs per SM of their grids.  Defaults leave a field unchanged;
    returns the mode in effect (csrc/norm_kernels.cu)."""
    return int(_lib().b200_ln_config(int(pipe), int(fwd_ctas), int(bwd_ctas)))


def ln_bwd(dy, x, gamma, beta, mean, rstd, dgamma, dbeta, *, bias=None, dbias=None, residual=None, eps=1e-12, p_pre=0.0,
           p_post=0.0, seed=None, stream_pre=0, stream_post=0, want_ds=True, want_dz=False, ws: Optional[Workspace] = None,
           deterministic: bool = False):
    """Returns (ds, dz): ds = grad wrt the LN input (residual branch), dz = grad wrt x (through pre-dropout).
    Parameter gradients are accumulated into dgamma / dbeta / dbias (fp32)."""
    ws = ws or Workspace.get(x.device)
    rows, cols = x.shape
    a = _ln_base(x, bias, residual, gamma, beta, eps, p_pre, p_post, seed, stream_pre, stream_post)
    a.mean, a.rstd = mean.data_ptr(), rstd.data_ptr()
    a.dy, a.lddy = dy.data_ptr(), dy.stride(0)
    ds = torch.empty((rows, cols), dtype=x.dtype, device=x.device) if want_ds else None
    dz = torch.empty((rows, cols), dtype=x.dtype, device=x.device) if want_dz else None
    if ds is not None:
        a.ds, a.ldds = ds.data_ptr(), ds.stride(0)
    if dz is not None:
        a.dz, a.lddz = dz.data_ptr(), dz.stride(0)
    a.dgamma, a.dbeta, a.dbias = dgamma.data_ptr(), dbeta.data_ptr(), _ptr(dbias)
    if deterministic or cols % 256 != 0 or _LN_TWO_PASS:
        sc = ws.ln_scratch(cols)
        a.scratch, a.scratch_bytes = sc.data_ptr(), sc.numel() * 4
    else:
        a.scratch, a.scratch_bytes = None, 0       # block partials are added to the parameter gradients directly
    check(_lib().b200_ln_bwd(byref(a), _stream()), "b200_ln_bwd")
    return ds, dz


# ---------------------------------------------------------------------------------------------------
# embeddings / compaction / losses / heads
# ---------------------------------------------------------------------------------------------------
def embed_sum(ids, tts, W, P, TT, L):
    rows = ids.numel()
    H = W.shape[1]
    out = torch.empty((rows, H), dtype=W.dtype, device=W.device)
    check(_lib().b200_embed_sum(dtype_code(W.dtype), ids.data_ptr(), _ptr(tts), W.data_ptr(), P.data_ptr(), TT.data_ptr(),
                                out.data_ptr(), rows, L, H, W.shape[0], TT.shape[0], _stream()), "b200_embed_sum")
    return out


def embed_scatter(ids, tts, ds, gW, gP, gT, L):
    rows, H = ds.shape
    check(_lib().b200_embed_scatter(dtype_code(ds.dtype), ids.data_ptr(), _ptr(tts), ds.data_ptr(), gW.data_ptr(),
                                    gP.data_ptr(), gT.data_ptr(), rows, L, H, gW.shape[0], gT.shape[0], _stream()),
          "b200_embed_scatter")


def compact_rows(labels, cap, ignore=-1, mask=None):
    rows = labels.numel()
    idx = torch.empty(cap, dtype=torch.int32, device=labels.device)
    lab = torch.empty(cap, dtype=torch.int64, device=labels.device)
    cnt = torch.empty(2, dtype=torch.int32, device=labels.device)
    check(_lib().b200_compact_rows(labels.data_ptr(), _ptr(mask), rows, ignore, cap, idx.data_ptr(), lab.data_ptr(),
                                   cnt.data_ptr(), _stream()), "b200_compact_rows")
    return idx, lab, cnt


def gather_rows(x, idx):
    n, H = idx.numel(), x.shape[1]
    out = torch.empty((n, H), dtype=x.dtype, device=x.device)
    check(_lib().b200_gather_rows(dtype_code(x.dtype), x.data_ptr(), x.stride(0), idx.data_ptr(), n, H, out.data_ptr(),
                                  _stream()), "b200_gather_rows")
    return out


def scatter_rows(dy, idx, dx, accumulate=False):
    n, H = dy.shape
    check(_lib().b200_scatter_rows(dtype_code(dy.dtype), dy.data_ptr(), idx.data_ptr(), n, H, dx.data_ptr(), dx.stride(0),
                                   int(accumulate), _stream()), "b200_scatter_rows")


def ce_fwd(logits, labels, ignore=-1):
    rows, V = logits.shape
    lse = torch.empty(rows, dtype=torch.float32, device=logits.device)
    row_loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
    check(_lib().b200_ce_fwd(dtype_code(logits.dtype), logits.data_ptr(), logits.stride(0), labels.data_ptr(), rows, V,
                             ignore, lse.data_ptr(), row_loss.data_ptr(), _stream()), "b200_ce_fwd")
    return lse, row_loss


def ce_bwd_(logits, labels, lse, gout=None, inv_count=None, ignore=-1):
    rows, V = logits.shape
    check(_lib().b200_ce_bwd(dtype_code(logits.dtype), logits.data_ptr(), logits.stride(0), labels.data_ptr(), rows, V,
                             ignore, lse.data_ptr(), _ptr(gout), _ptr(inv_count), _stream()), "b200_ce_bwd")
    return logits


def ce_finalize(row_loss, labels, ignore=-1):
    """Returns (loss[1], inv_count[1], count[1]) device tensors."""
    dev = row_loss.device
    out = torch.empty(3, dtype=torch.float32, device=dev)
    check(_lib().b200_ce_finalize(row_loss.data_ptr(), labels.data_ptr(), row_loss.numel(), ignore, out[0:1].data_ptr(),
                                  out[1:2].data_ptr(), out[2:3].data_ptr(), _stream()), "b200_ce_finalize")
    return out[0:1], out[1:2], out[2:3]


def head_fwd(x, w, b):
    rows, K = x.shape
    N = w.shape[0]
    out = torch.empty((rows, N), dtype=torch.float32, device=x.device)
    check(_lib().b200_head_fwd(dtype_code(x.dtype), x.data_ptr(), x.stride(0), w.data_ptr(), _ptr(b), rows, K, N,
                               out.data_ptr(), _stream()), "b200_head_fwd")
    return out


def head_ce(logits, labels, active=None, ignore=-100):
    rows, N = logits.shape
    row_loss = torch.empty(rows, dtype=torch.float32, device=logits.device)
    eff = torch.empty(rows, dtype=torch.int64, device=logits.device)
    check(_lib().b200_head_ce(logits.data_ptr(), labels.data_ptr(), _ptr(active), rows, N, ignore, row_loss.data_ptr(),
                              eff.data_ptr(), _stream()), "b200_head_ce")
    return row_loss, eff


def head_ce_bwd(logits, eff_labels, gout=None, inv_count=None, ignore=-100):
    rows, N = logits.shape
    dl = torch.empty_like(logits)
    check(_lib().b200_head_ce_bwd(logits.data_ptr(), eff_labels.data_ptr(), rows, N, ignore, _ptr(gout), _ptr(inv_count),
                                  dl.data_ptr(), _stream()), "b200_head_ce_bwd")
    return dl


def head_bwd(dl, x, w, dx=None, accumulate_dx=False, dw=None, db=None):
    rows, K = x.shape
    N = w.shape[0]
    check(_lib().b200_head_bwd(dtype_code(x.dtype), dl.data_ptr(), x.data_ptr(), x.stride(0), w.data_ptr(), rows, K, N,
                               _ptr(dx), dx.stride(0) if dx is not None else 0, int(accumulate_dx), _ptr(dw), _ptr(db),
                               _stream()), "b200_head_bwd")


def colsum_(x, out, ws: Optional[Workspace] = None, deterministic: bool = False):
    """out[c] += sum_r x[r, c]  (fp32 accumulation).  Default: single pass with fp32 reductions into ``out``;
    ``deterministic=True``: two passes through workspace scratch with a fixed summation order."""
    rows, cols = x.shape
    if deterministic or (out.data_ptr() & 15):
        ws = ws or Workspace.get(x.device)
        sc = ws.colsum_scratch(rows, cols)
        check(_lib().b200_colsum(dtype_code(x.dtype), x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), sc.data_ptr(),
                                 sc.numel() * 4, _stream()), "b200_colsum")
    else:
        check(_lib().b200_colsum(dtype_code(x.dtype), x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), None, 0,
                                 _stream()), "b200_colsum")
    return out


class SideWork:
    """Memory-bound side computations that nothing downstream in the layer waits for (bias-gradient column sums) can run
    on a second stream next to the tensor-core kernels of the main stream, joined before the layer announces its
    gradients; inside a captured CUDA graph this becomes a parallel branch.  Measured on B200 (profiles/README.md §18):
    the 24 fork / join pairs per step cost more than the ~10 us kernels they hide (7.83 vs 7.70 ms / step), so it is
    opt-in: B200_SIDE_COLSUM=1."""

    _per_device = {}

    def __init__(self, device):
        self.device = device
        self.stream = torch.cuda.Stream(device=device)
        self.enabled = bool(os.environ.get("B200_SIDE_COLSUM"))
        self._dirty = False
        self._keep = []

    @classmethod
    def get(cls, device) -> "SideWork":
        device = torch.device(device)
        idx = device.index if device.index is not None else torch.cuda.current_device()
        sw = cls._per_device.get(idx)
        if sw is None:
            sw = cls._per_device[idx] = SideWork(torch.device("cuda", idx))
        return sw

    def colsum_(self, x, out):
        """out[c] += sum_r x[r, c] on the side stream (x must stay unchanged until ``join``)."""
        if not self.enabled or (out.data_ptr() & 15):
            return colsum_(x, out)
        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)
        self.stream.wait_event(ev)
        with torch.cuda.stream(self.stream):
            colsum_(x, out)
        self._keep.append(x)               # its memory must not be recycled before the side stream is done with it
        self._dirty = True
        return out

    def join(self):
        """Make the current stream wait for everything issued on the side stream."""
        if self._dirty:
            torch.cuda.current_stream(self.device).wait_stream(self.stream)
            self._dirty = False
            self._keep.clear()


def softmax_fwd_(S, attn_mask, B, heads, L, Lp, p_drop=0.0, seed=None, stream_id=0):
    """In place P = softmax(S + mask); returns Pdrop (a new tensor) when p_drop > 0 else S itself."""
    pdrop = torch.empty_like(S) if p_drop > 0 else None
    check(_lib().b200_softmax_fwd(dtype_code(S.dtype), S.data_ptr(), _ptr(pdrop), _ptr(attn_mask), B, heads, L, Lp,
                                  float(p_drop), _ptr(seed), stream_id, _stream()), "b200_softmax_fwd")
    return pdrop if pdrop is not None else S


def softmax_bwd_(P, dP, B, heads, L, Lp, p_drop=0.0, seed=None, stream_id=0):
    check(_lib().b200_softmax_bwd(dtype_code(P.dtype), P.data_ptr(), dP.data_ptr(), B, heads, L, Lp, float(p_drop),
                                  _ptr(seed), stream_id, _stream()), "b200_softmax_bwd")
    return dP


def assemble_bert_batch(raw, L, P, out=None):
    """Packed int32 batch [B, 3L+2P+1] (device) -> [input_ids, segment_ids, input_mask, masked_lm_labels,
    next_sentence_labels] as int64 device tensors (ref: hetseq/data/h5pyDataset.py:33-48)."""
    assert raw.dtype == torch.int32 and raw.is_cuda and raw.is_contiguous() and raw.shape[1] == 3 * L + 2 * P + 1
    B = raw.shape[0]
    if out is None:
        out = [torch.empty((B, L), dtype=torch.int64, device=raw.device) for _ in range(4)]
        out.append(torch.empty((B,), dtype=torch.int64, device=raw.device))
    ids, seg, mask, labels, nsl = out
    check(_lib().b200_assemble_bert_batch(raw.data_ptr(), B, L, P, ids.data_ptr(), seg.data_ptr(), mask.data_ptr(),
                                          labels.data_ptr(), nsl.data_ptr(), _stream()), "b200_assemble_bert_batch")
    return out


def _attn_args(qkv, attn_mask, B, heads, L, Lp, scale, p_drop, seed, stream_id):
    assert qkv.dim() == 2 and qkv.shape[0] == B * L and qkv.shape[1] == 3 * heads * 64 and qkv.stride(1) == 1
    a = AttnArgs()
    a.dtype = dtype_code(qkv.dtype)
    a.B, a.heads, a.L, a.Lp = B, heads, L, Lp
    a.scale, a.p_drop = float(scale), float(p_drop)
    a.qkv, a.ld_qkv = qkv.data_ptr(), qkv.stride(0)
    a.attn_mask = _ptr(attn_mask)
    a.seed_dev = _ptr(seed)
    a.stream_id = stream_id
    return a


def attn_fwd(qkv, attn_mask, B, heads, L, scale, p_drop=0.0, seed=None, stream_id=0, Lp=None, out=None):
    """Fused attention forward on the packed projection ``qkv`` [B*L, 3*heads*64].

    Returns (ctx [B*L, heads*64], lse [B*heads*L] fp32).  ref: hetseq/bert_modeling.py:341-376."""
    Lp = Lp or (L + 3) // 4 * 4
    a = _attn_args(qkv, attn_mask, B, heads, L, Lp, scale, p_drop, seed, stream_id)
    if out is None:
        out = torch.empty(B * L, heads * 64, dtype=qkv.dtype, device=qkv.device)
    lse = torch.empty(B * heads * L, dtype=torch.float32, device=qkv.device)
    a.out, a.ld_out, a.lse = out.data_ptr(), out.stride(0), lse.data_ptr()
    check(_lib().b200_attn_fwd(byref(a), _stream()), "b200_attn_fwd")
    return out, lse


def attn_bwd(qkv, o, dout, lse, attn_mask, B, heads, L, scale, p_drop=0.0, seed=None, stream_id=0, Lp=None, out=None):
    """Fused attention backward: returns dqkv [B*L, 3*heads*64] (every element written)."""
    Lp = Lp or (L + 3) // 4 * 4
    a = _attn_args(qkv, attn_mask, B, heads, L, Lp, scale, p_drop, seed, stream_id)
    assert o.stride(1) == 1 and dout.stride(1) == 1 and o.stride(0) == dout.stride(0) and o.dtype == dout.dtype == qkv.dtype
    if out is None:
        out = torch.empty_like(qkv)
    delta = torch.empty(B * heads * L, dtype=torch.float32, device=qkv.device)
    a.out, a.ld_out, a.lse, a.delta = out.data_ptr(), out.stride(0), lse.data_ptr(), delta.data_ptr()
    a.o, a.dout, a.ld_o = o.data_ptr(), dout.data_ptr(), o.stride(0)
    check(_lib().b200_attn_bwd(byref(a), _stream()), "b200_attn_bwd")
    return out


def dropout(x, p, seed, stream_id, out=None):
    """y = x * keep_mask / (1-p); calling it again on a gradient with the same ids is the backward."""
    if out is None:
        out = torch.empty_like(x)
    assert x.is_contiguous() and out.is_contiguous()
    check(_lib().b200_dropout(dtype_code(x.dtype), x.data_ptr(), out.data_ptr(), x.numel(), float(p), _ptr(seed), stream_id,
                              _stream()), "b200_dropout")
    return out


def act_bwd(dy, aux, mode: int):
    """dy * act'(aux): mode 1 = gelu' at the pre-activation, mode 2 = tanh' at the tanh output."""
    dy = dy.contiguous()
    aux = aux.contiguous()
    out = torch.empty_like(dy)
    check(_lib().b200_act_bwd(dtype_code(dy.dtype), dy.data_ptr(), aux.data_ptr(), out.data_ptr(), dy.numel(), mode,
                              _stream()), "b200_act_bwd")
    return out
