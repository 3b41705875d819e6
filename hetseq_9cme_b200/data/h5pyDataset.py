"""Pre-tokenised BERT shard datasets (API of hetseq/data/h5pyDataset.py:13-134).

``BertH5pyData`` keeps the reference's name and sample contract — ``dataset[i]`` is
``[input_ids, segment_ids, input_mask, masked_lm_labels, next_sentence_labels]`` (int64), the positional
argument order of ``BertForPreTraining.forward`` — but opens the shard ONCE (memory-mapped ``.b2s`` or a fully
read HDF5 file) and can assemble a whole batch with vectorised NumPy slicing (``fetch_batch``) instead of six
h5py row reads and a file open per sample.
"""
from __future__ import annotations

import ctypes
from typing import List, Sequence

import numpy as np
import torch
import torch.utils.data

from ..ops import native
from . import shard_io


def _scatter_labels(positions: np.ndarray, ids: np.ndarray, L: int) -> np.ndarray:
    """Dense labels: -1 everywhere except ``labels[r, positions[r, :k]] = ids[r, :k]`` where k = index of the
    first zero in ``positions[r]`` (ref: hetseq/data/h5pyDataset.py:42-48)."""
    positions = np.ascontiguousarray(positions, dtype=np.int64)
    ids = np.ascontiguousarray(ids, dtype=np.int64)
    rows, P = positions.shape
    out = np.empty((rows, L), dtype=np.int64)
    lib = native.host_lib()
    if lib is not None:
        fn = lib.b200_scatter_mlm_labels
        fn.restype = None
        fn.argtypes = [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_longlong, ctypes.c_void_p]
        fn(positions.ctypes.data, ids.ctypes.data, rows, P, L, out.ctypes.data)
        return out
    out.fill(-1)
    valid = np.cumprod(positions != 0, axis=1).astype(bool)
    r, c = np.nonzero(valid)
    out[r, positions[r, c]] = ids[r, c]
    return out


class BertH5pyData(torch.utils.data.Dataset):
    keys = shard_io.KEYS

    def __init__(self, path, max_pred_length=512):
        super().__init__()
        self.path = path
        self.max_pred_length = max_pred_length
        self._shard = shard_io.open_shard(path) if isinstance(path, str) else path
        self._len = len(self._shard[self.keys[0]])
        self.seq_len = int(self._shard[self.keys[0]].shape[1])

    def __len__(self):
        return self._len

    def check_index(self, i):
        if i < 0 or i >= self._len:
            raise IndexError("index out of range")

    def fetch_rows(self, rows: Sequence[int]) -> List[torch.Tensor]:
        rows = np.asarray(rows, dtype=np.int64)
        sh = self._shard
        take = lambda k: np.asarray(sh[k][rows] if not _is_sorted_range(rows) else sh[k][rows[0]:rows[-1] + 1]).astype(np.int64, copy=False)
        input_ids, input_mask, segment_ids = take("input_ids"), take("input_mask"), take("segment_ids")
        labels = _scatter_labels(take("masked_lm_positions"), take("masked_lm_ids"), input_ids.shape[1])
        nsl = take("next_sentence_labels").reshape(-1)
        return [torch.from_numpy(np.ascontiguousarray(a)) for a in (input_ids, segment_ids, input_mask, labels, nsl)]

    def fetch_rows_packed(self, rows: Sequence[int]) -> np.ndarray:
        """The same rows as one packed int32 matrix for the device-side assembly (data/packed.py)."""
        from .packed import pack_rows
        rows = np.asarray(rows, dtype=np.int64)
        sh = self._shard
        take = lambda k: np.asarray(sh[k][rows] if not _is_sorted_range(rows) else sh[k][rows[0]:rows[-1] + 1])
        return pack_rows(take("input_ids"), take("segment_ids"), take("input_mask"), take("masked_lm_positions"),
                         take("masked_lm_ids"), take("next_sentence_labels"))

    @property
    def max_pred(self) -> int:
        return int(self._shard["masked_lm_positions"].shape[1])

    def __getitem__(self, index):
        self.check_index(index)
        ids, seg, mask, lab, nsl = self.fetch_rows([index])
        return [ids[0], seg[0], mask[0], lab[0], nsl[0]]

    def size(self, idx: int):
        return self.max_pred_length

    def set_epoch(self, epoch):
        pass


def _distinct(a: np.ndarray) -> List[int]:
    """Sorted distinct values.  Not ``np.unique``: its first call imports ``numpy.ma`` (~50 ms under the GIL), which the
    loader thread would spend in the middle of a training step — measured as a 20-40 ms GPU bubble at the first
    shard-spanning batch (profiles/README.md §8)."""
    return sorted(set(a.tolist()))


def _is_sorted_range(rows: np.ndarray) -> bool:
    return len(rows) > 0 and rows[-1] - rows[0] == len(rows) - 1 and (len(rows) == 1 or bool(np.all(np.diff(rows) == 1)))


class ConBertH5pyData(torch.utils.data.Dataset):
    """Concatenation of shard datasets addressed by one global index (cumulative sizes + binary search)."""

    @staticmethod
    def cumsum(sequence, sample_ratios):
        out, s = [], 0
        for e, ratio in zip(sequence, sample_ratios):
            s += int(ratio * len(e))
            out.append(s)
        return out

    def __init__(self, datasets, sample_ratios=1):
        super().__init__()
        self.datasets = list(datasets)
        assert len(self.datasets) > 0, "datasets should not be an empty iterable"
        if isinstance(sample_ratios, int):
            sample_ratios = [sample_ratios] * len(self.datasets)
        self.sample_ratios = sample_ratios
        self.cumulative_sizes = self.cumsum(self.datasets, sample_ratios)
        self.real_sizes = [len(d) for d in self.datasets]
        self._cum = np.asarray(self.cumulative_sizes, dtype=np.int64)

    def __len__(self):
        return self.cumulative_sizes[-1]

    def _locate(self, idx: np.ndarray):
        ds = np.searchsorted(self._cum, idx, side="right")
        local = idx - np.where(ds > 0, self._cum[np.maximum(ds - 1, 0)], 0)
        local = local % np.asarray(self.real_sizes, dtype=np.int64)[ds]
        return ds, local

    def _get_dataset_and_sample_index(self, idx: int):
        ds, local = self._locate(np.asarray([idx], dtype=np.int64))
        return int(ds[0]), int(local[0])

    def __getitem__(self, idx):
        d, s = self._get_dataset_and_sample_index(idx)
        return self.datasets[d][s]

    def fetch_batch(self, indices: Sequence[int]):
        """Collated batch for ``indices`` without touching individual samples."""
        idx = np.asarray(indices, dtype=np.int64)
        ds, local = self._locate(idx)
        if not hasattr(self.datasets[0], "fetch_rows"):
            return self.collater([self[int(i)] for i in idx])
        if getattr(self, "packed", False) and hasattr(self.datasets[0], "fetch_rows_packed"):
            from .packed import PackedBertBatch
            raw = np.empty((len(idx), 0), dtype=np.int32)
            if np.all(ds == ds[0]):
                raw = self.datasets[int(ds[0])].fetch_rows_packed(local)
            else:
                first = self.datasets[int(ds[0])]
                raw = np.empty((len(idx), 3 * first.seq_len + 2 * first.max_pred + 1), dtype=np.int32)
                for d in _distinct(ds):
                    raw[ds == d] = self.datasets[int(d)].fetch_rows_packed(local[ds == d])
            first = self.datasets[int(ds[0])]
            return PackedBertBatch.build(raw, first.seq_len, first.max_pred)
        if np.all(ds == ds[0]):
            return self.datasets[int(ds[0])].fetch_rows(local)
        parts = [self.datasets[int(d)].fetch_rows(local[ds == d]) for d in _distinct(ds)]
        order = np.argsort(np.concatenate([np.nonzero(ds == d)[0] for d in _distinct(ds)]))
        return [torch.cat([p[k] for p in parts])[order] for k in range(len(parts[0]))]

    def collater(self, samples):
        if len(samples) == 0:
            return None
        if hasattr(self.datasets[0], "collater"):
            return self.datasets[0].collater(samples)
        return torch.utils.data.default_collate(samples)

    def ordered_indices(self):
        return np.arange(len(self))

    def sizes_array(self) -> np.ndarray:
        """Per-sample token counts for the batcher (constant per shard)."""
        out = np.empty(len(self), dtype=np.int64)
        start = 0
        for d, end in zip(self.datasets, self.cumulative_sizes):
            out[start:end] = int(np.max(d.size(0)))
            start = end
        return out

    def num_tokens(self, index: int):
        return np.max(self.size(index))

    def size(self, idx: int):
        d, s = self._get_dataset_and_sample_index(idx)
        return self.datasets[d].size(s)

    def set_epoch(self, epoch):
        pass
