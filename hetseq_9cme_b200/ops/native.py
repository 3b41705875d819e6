"""ctypes binding of the in-tree native libraries.

``libb200seq.so`` holds every hand-written sm_100a kernel (C ABI declared in ``include/b200seq.h``);
``libb200seq_host.so`` holds the host-only C++ runtime pieces (batcher, shard index).  Both are built
in-tree by ``tools/build_native.py`` / ``__graft_entry__.build()``.

The CUDA library is *required* whenever a CUDA device is present: there is no silent eager fallback.
``cuda_lib()`` raises if the shared object is missing, so a GPU run can never quietly execute a
PyTorch-only path.
"""
from __future__ import annotations

import ctypes
import os
import threading
from pathlib import Path

_LIB_DIR = Path(__file__).resolve().parent.parent / "lib"
_lock = threading.Lock()
_cuda = None
_host = None

c_void_p = ctypes.c_void_p
c_int = ctypes.c_int
c_ll = ctypes.c_longlong
c_float = ctypes.c_float


class NativeError(RuntimeError):
    pass


# ---- enums (mirror include/b200seq.h) -------------------------------------------------------------
F32_TF32, BF16, BF16_F32OUT = 0, 1, 2
ACT_NONE, ACT_GELU, ACT_TANH, ACT_GELU_SAVE_GRAD = 0, 1, 2, 3
AUX_NONE, AUX_ADD, AUX_MUL_DGELU, AUX_MUL_DTANH, AUX_MUL = 0, 1, 2, 3, 4


class GemmDesc(ctypes.Structure):
    _fields_ = [
        ("dtype", c_int), ("M", c_int), ("N", c_int), ("K", c_int),
        ("A", c_void_p), ("lda", c_ll), ("a_mn", c_int),
        ("B", c_void_p), ("ldb", c_ll), ("b_mn", c_int),
        ("D", c_void_p), ("ldd", c_ll),
        ("D2", c_void_p),
        ("bias", c_void_p),
        ("aux", c_void_p), ("ldaux", c_ll),
        ("act", c_int), ("aux_mode", c_int),
        ("accumulate", c_int),
        ("alpha", c_float),
        ("bn", c_int), ("splits", c_int),
        ("batch0", c_int), ("batch1", c_int),
        ("sa0", c_ll), ("sa1", c_ll), ("sb0", c_ll), ("sb1", c_ll), ("sd0", c_ll), ("sd1", c_ll),
        ("epi_mode", c_int), ("attn_mask", c_void_p), ("p_drop", c_float), ("seed_dev", c_void_p),
        ("stream_id", ctypes.c_ulonglong),
    ]


def lib_path(name: str = "libb200seq.so") -> Path:
    # B200SEQ_LIB: load another build of the CUDA library (A/B runs against an older build of the kernels)
    override = os.environ.get("B200SEQ_LIB")
    if override and name == "libb200seq.so":
        return Path(override)
    return _LIB_DIR / name


def cuda_lib_available() -> bool:
    return lib_path().exists()


def _declare(lib):
    lib.b200_last_error.restype = ctypes.c_char_p
    lib.b200_version.restype = c_int
    lib.b200_launch_count.restype = c_ll
    lib.b200_reset_launch_count.restype = None
    lib.b200_add_launch_count.argtypes = [c_ll]
    lib.b200_add_launch_count.restype = None
    lib.b200_gemm.argtypes = [ctypes.POINTER(GemmDesc), c_void_p]
    lib.b200_gemm.restype = c_int
    # every other entry point takes (pointers/ints/floats..., stream) and returns int; argtypes are
    # declared where the op wrappers live (ops/kernels.py) to keep this module small.


def cuda_lib():
    """Return the loaded ``libb200seq.so`` (loads on first use; raises NativeError if absent)."""
    global _cuda
    if _cuda is None:
        with _lock:
            if _cuda is None:
                p = lib_path()
                if not p.exists():
                    raise NativeError(
                        f"{p} is missing: run `python tools/build_native.py` (or __graft_entry__.build()) — "
                        "the CUDA path has no PyTorch fallback on a GPU machine")
                lib = ctypes.CDLL(str(p), mode=os.RTLD_GLOBAL if hasattr(os, "RTLD_GLOBAL") else ctypes.DEFAULT_MODE)
                _declare(lib)
                _cuda = lib
    return _cuda


def host_lib():
    """Return the loaded host-only library, or None when it has not been built."""
    global _host
    if _host is None:
        with _lock:
            if _host is None:
                p = lib_path("libb200seq_host.so")
                if p.exists():
                    _host = ctypes.CDLL(str(p))
    return _host


def check(rc: int, what: str = "native call"):
    if rc != 0:
        msg = cuda_lib().b200_last_error()
        raise NativeError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")


def launch_count() -> int:
    return int(cuda_lib().b200_launch_count()) if _cuda is not None or cuda_lib_available() else 0


def add_launch_count(n: int) -> None:
    cuda_lib().b200_add_launch_count(int(n))


def reset_launch_count() -> None:
    if _cuda is not None or cuda_lib_available():
        cuda_lib().b200_reset_launch_count()
