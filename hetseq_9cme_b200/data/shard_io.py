"""Shard storage for pre-tokenised BERT corpora.

The reference reads NVIDIA-style HDF5 shards through h5py, re-opening the file for every sample
(hetseq/data/h5pyDataset.py:31-51).  h5py / libhdf5 are not part of this image, so shards are handled by

* ``B2Shard``  — our own flat container: a JSON header followed by 64-byte-aligned raw arrays, opened once
  and memory-mapped (zero-copy row slicing, shareable across ranks through the page cache);
* ``open_shard(path)`` — dispatches on the file: ``.b2s`` -> B2Shard, ``.hdf5/.h5`` -> h5py when importable,
  else the minimal built-in HDF5 reader (``hdf5_lite``), loading each dataset once.

Both expose ``shard[key] -> np.ndarray`` with the six keys of the reference schema.
"""
from __future__ import annotations

import json
import os
import struct
from typing import Dict, Mapping

import numpy as np

MAGIC = b"B2SHARD1"
KEYS = ("input_ids", "input_mask", "segment_ids", "masked_lm_positions", "masked_lm_ids", "next_sentence_labels")


def write_b2shard(path: str, arrays: Mapping[str, np.ndarray]) -> None:
    """Atomically write ``arrays`` as one ``.b2s`` file."""
    entries = []
    offset = 0
    blobs = []
    for name, arr in arrays.items():
        arr = np.ascontiguousarray(arr)
        pad = (-offset) % 64
        offset += pad
        entries.append({"name": name, "dtype": arr.dtype.str, "shape": list(arr.shape), "offset": offset, "nbytes": arr.nbytes})
        blobs.append((pad, arr))
        offset += arr.nbytes
    header = json.dumps({"arrays": entries}).encode("utf-8")
    hpad = (-(len(MAGIC) + 8 + len(header))) % 64
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<Q", len(header) + hpad))
        f.write(header + b" " * hpad)
        for pad, arr in blobs:
            f.write(b"\0" * pad)
            f.write(arr.tobytes())
    os.replace(tmp, path)


class B2Shard:
    """Read-only, memory-mapped view of a ``.b2s`` file."""

    def __init__(self, path: str):
        self.path = path
        with open(path, "rb") as f:
            if f.read(len(MAGIC)) != MAGIC:
                raise ValueError("{} is not a b2shard file".format(path))
            (hlen,) = struct.unpack("<Q", f.read(8))
            header = json.loads(f.read(hlen).decode("utf-8"))
        base = len(MAGIC) + 8 + hlen
        self._arrays: Dict[str, np.ndarray] = {}
        for e in header["arrays"]:
            self._arrays[e["name"]] = np.memmap(path, mode="r", dtype=np.dtype(e["dtype"]), shape=tuple(e["shape"]),
                                                offset=base + e["offset"])

    def __getitem__(self, key):
        return self._arrays[key]

    def keys(self):
        return self._arrays.keys()

    def __contains__(self, key):
        return key in self._arrays


class _DictShard:
    def __init__(self, arrays):
        self._arrays = arrays

    def __getitem__(self, key):
        return self._arrays[key]

    def keys(self):
        return self._arrays.keys()

    def __contains__(self, key):
        return key in self._arrays


def open_shard(path: str):
    low = path.lower()
    if low.endswith(".b2s"):
        return B2Shard(path)
    if low.endswith((".hdf5", ".h5", ".hdf")):
        try:
            import h5py  # type: ignore
            if not hasattr(h5py, "File") or not hasattr(h5py, "version"):
                raise ImportError("h5py stub")
            with h5py.File(path, "r") as f:
                return _DictShard({k: np.asarray(f[k][:]) for k in f.keys()})
        except ImportError:
            from .hdf5_lite import read_hdf5
            return _DictShard(read_hdf5(path))
    if low.endswith(".npz"):
        with np.load(path) as z:
            return _DictShard({k: z[k] for k in z.files})
    raise ValueError("unknown shard format: {}".format(path))


def convert_to_b2shard(src: str, dst: str) -> None:
    sh = open_shard(src)
    write_b2shard(dst, {k: np.asarray(sh[k]) for k in sh.keys()})
