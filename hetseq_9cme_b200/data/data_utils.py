"""Batch construction helpers (API parity with hetseq/data/data_utils.py:14-61).

``batch_by_size`` accepts either a sizes array or the reference's ``num_tokens_fn`` callback; the work is done
by the C++ routine in ``csrc/host/batcher.cpp`` (a NumPy restatement is used if the host library is missing).
"""
from __future__ import annotations

import contextlib
import ctypes
import sys
from typing import Callable, List, Optional, Sequence, Union

import numpy as np

from ..ops import native


@contextlib.contextmanager
def numpy_seed(seed, *addl_seeds):
    """Seed NumPy's global generator inside the block and restore its state afterwards."""
    if seed is None:
        yield
        return
    if addl_seeds:
        seed = int(hash((seed, *addl_seeds)) % 1e6)
    state = np.random.get_state()
    np.random.seed(seed)
    try:
        yield
    finally:
        np.random.set_state(state)


def _cuts_python(sizes: np.ndarray, max_tokens: int, max_sentences: int, bsz_mult: int) -> List[int]:
    cuts = [0]
    start, sample_len = 0, 0
    n = len(sizes)
    for i in range(n):
        sample_len = max(sample_len, int(sizes[i]))
        if sample_len > max_tokens:
            raise AssertionError("sentence at position {} of size {} exceeds max_tokens limit of {}!".format(i, sample_len, max_tokens))
        ln = i - start
        if ln > 0 and (ln == max_sentences or (ln + 1) * sample_len > max_tokens):
            mod_len = max(bsz_mult * (ln // bsz_mult), ln % bsz_mult)
            start += mod_len
            cuts.append(start)
            sample_len = int(sizes[start:i + 1].max())
    if start < n:
        cuts.append(n)
    return cuts


def batch_cuts(sizes: np.ndarray, max_tokens: Optional[int] = None, max_sentences: Optional[int] = None,
               required_batch_size_multiple: int = 1) -> np.ndarray:
    """Boundaries of the greedy batches over an ordered list whose element sizes are ``sizes``."""
    sizes = np.ascontiguousarray(sizes, dtype=np.int64)
    mt = int(max_tokens) if max_tokens is not None else sys.maxsize
    ms = int(max_sentences) if max_sentences is not None else sys.maxsize
    lib = native.host_lib()
    if lib is None:
        return np.asarray(_cuts_python(sizes, mt, ms, required_batch_size_multiple), dtype=np.int64)
    fn = lib.b200_batch_by_size
    fn.restype = ctypes.c_longlong
    fn.argtypes = [ctypes.c_void_p, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_void_p]
    cuts = np.empty(len(sizes) + 1, dtype=np.int64)
    nb = fn(sizes.ctypes.data, len(sizes), mt, ms, required_batch_size_multiple, cuts.ctypes.data)
    if nb < 0:
        i = -nb - 1
        raise AssertionError("sentence at position {} of size {} exceeds max_tokens limit of {}!".format(i, int(sizes[i]), mt))
    return cuts[: nb + 1]


def batch_by_size(indices: Sequence[int], num_tokens_fn: Union[Callable[[int], int], np.ndarray, None] = None,
                  max_tokens: Optional[int] = None, max_sentences: Optional[int] = None,
                  required_batch_size_multiple: int = 1, sizes: Optional[np.ndarray] = None) -> List[List[int]]:
    """Mini-batches of indices bucketed by size (same cut rules as the reference).

    ``sizes`` (or an array passed as ``num_tokens_fn``) is indexed by dataset index; a callable is evaluated
    once per element for compatibility with reference-style datasets."""
    indices = np.asarray(indices, dtype=np.int64)
    if sizes is None and isinstance(num_tokens_fn, np.ndarray):
        sizes = num_tokens_fn
    if sizes is not None:
        sz = np.asarray(sizes, dtype=np.int64)
        per_elem = sz[indices] if sz.ndim == 1 and len(sz) > 0 else np.full(len(indices), int(sz), dtype=np.int64)
    elif num_tokens_fn is not None:
        per_elem = np.fromiter((num_tokens_fn(int(i)) for i in indices), dtype=np.int64, count=len(indices))
    else:
        per_elem = np.ones(len(indices), dtype=np.int64)
    cuts = batch_cuts(per_elem, max_tokens, max_sentences, required_batch_size_multiple)
    return [indices[cuts[b]:cuts[b + 1]].tolist() for b in range(len(cuts) - 1)]
