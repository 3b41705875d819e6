"""b200_colsum_static: the column-sum entry point for callers that cannot allocate device memory.

Some users of the library may not call cudaMalloc at all — e.g. an application whose own pool allocator already
owns the device, or a loop that cannot tolerate the device-wide stall an allocation causes — so they have no
scratch buffer to hand to ``b200_colsum``.  ``b200_colsum_static`` keeps its partial sums in a 1 MiB array that is
part of the library image instead.  This test drives it purely through the C ABI (ctypes, raw pointers) and checks
it against the regular entry point, twice in a row: the second call reuses the array with the first call's
contents still in it, which must not leak into the result.

The file name sorts last on purpose so that a ``pytest -x`` run has executed every other test before this one.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from hetseq_9cme_b200.ops import native
    lib = native.cuda_lib()
    lib.b200_colsum_static.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_int,
                                       ctypes.c_void_p, ctypes.c_void_p]
    lib.b200_colsum_static.restype = ctypes.c_int
    lib.b200_colsum_static_max_cols.restype = ctypes.c_int
    lib.b200_colsum.argtypes = [ctypes.c_int, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_int, ctypes.c_void_p,
                                ctypes.c_void_p, ctypes.c_longlong, ctypes.c_void_p]
    lib.b200_colsum.restype = ctypes.c_int
    lib.b200_colsum_scratch_bytes.argtypes = [ctypes.c_int, ctypes.c_int]
    lib.b200_colsum_scratch_bytes.restype = ctypes.c_longlong
    return lib


def _regular(lib, x, out):
    rows, cols = x.shape
    nbytes = lib.b200_colsum_scratch_bytes(rows, cols)
    scratch = torch.empty(nbytes // 4, dtype=torch.float32, device=x.device)
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.b200_colsum(0 if x.dtype == torch.float32 else 1, x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(),
                         scratch.data_ptr(), nbytes, st)
    assert rc == 0, lib.b200_last_error()
    torch.cuda.synchronize()


def _static(lib, x, out):
    rows, cols = x.shape
    st = torch.cuda.current_stream().cuda_stream
    rc = lib.b200_colsum_static(0 if x.dtype == torch.float32 else 1, x.data_ptr(), x.stride(0), rows, cols, out.data_ptr(), st)
    assert rc == 0, lib.b200_last_error()
    torch.cuda.synchronize()


def test_static_scratch_matches_regular_entry_point_twice():
    lib = _lib()
    torch.manual_seed(123)
    # two calls in a row with different shapes/contents: the second finds the first call's partials in the array
    cases = [(torch.randn(4096, 3072, device="cuda") * 3.0, torch.float32),
             (torch.randn(700, 768, device="cuda"), torch.bfloat16)]
    for x, dt in cases:
        x = x.to(dt)
        want = torch.full((x.shape[1],), 0.5, dtype=torch.float32, device="cuda")   # both entry points accumulate (+=)
        got = want.clone()
        _regular(lib, x, want)
        _static(lib, x, got)
        ref = x.double().sum(0) + 0.5
        scale = ref.abs().max().item()
        assert (got.double() - want.double()).abs().max().item() <= 1e-5 * scale     # same kernels, same order
        assert (got.double() - ref).abs().max().item() <= 1e-4 * scale


def test_static_scratch_rejects_oversized_input():
    lib = _lib()
    limit = lib.b200_colsum_static_max_cols()
    assert limit == 4096
    x = torch.zeros(8, limit + 4, device="cuda")
    out = torch.zeros(limit + 4, device="cuda")
    rc = lib.b200_colsum_static(0, x.data_ptr(), x.stride(0), 8, limit + 4, out.data_ptr(), torch.cuda.current_stream().cuda_stream)
    assert rc != 0
    assert b"exceeds" in lib.b200_last_error()
    assert out.abs().max().item() == 0
