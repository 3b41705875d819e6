"""Packed BERT batches and their assembly on the device.

The reference builds every sample on the host: six h5py row reads, int64 casts and a Python scatter of
``masked_lm_ids`` into a dense ``masked_lm_labels`` row (hetseq/data/h5pyDataset.py:33-48), then the default
collate stacks them and ``utils.move_to_cuda`` copies five int64 tensors synchronously (hetseq/utils.py:32-37).

With ``--device-batch-assembly`` the loader thread only concatenates the shard's int32 rows into ONE matrix

    input_ids[L] | segment_ids[L] | input_mask[L] | masked_lm_positions[P] | masked_lm_ids[P] | next_sentence_label

pins it, and the controller copies it on a copy stream and lets ``b200_assemble_bert_batch``
(csrc/batch_kernels.cu) produce the five int64 tensors on the GPU: 4*(3L+2P+1) instead of 8*(4L+1) bytes per
sample over PCIe and no per-sample host work.
"""
from __future__ import annotations

from typing import List, Optional

import numpy as np
import torch


class PackedBertBatch(list):
    """``[raw int32 [B, 3L+2P+1], meta int32 [2] = (L, P)]`` — a list so that the generic sample helpers
    (pinning, emptiness test, shape signature) treat it like any other batch."""

    @classmethod
    def build(cls, raw: np.ndarray, seq_len: int, max_pred: int) -> "PackedBertBatch":
        assert raw.dtype == np.int32 and raw.ndim == 2 and raw.shape[1] == 3 * seq_len + 2 * max_pred + 1
        return cls([torch.from_numpy(np.ascontiguousarray(raw)), torch.tensor([seq_len, max_pred], dtype=torch.int32)])

    @property
    def raw(self) -> torch.Tensor:
        return self[0]

    @property
    def dims(self):
        L, P = (int(v) for v in self[1].tolist())      # meta stays on the host: no device sync
        return L, P

    def __len__(self):                                  # an empty batch must look empty to sample_is_empty()
        return 0 if list.__len__(self) == 0 or self[0].shape[0] == 0 else list.__len__(self)


def pack_rows(input_ids, segment_ids, input_mask, positions, ids, nsl) -> np.ndarray:
    """Concatenate the six shard arrays (any integer dtype) of n samples into the packed int32 matrix."""
    n = input_ids.shape[0]
    cols = [np.asarray(a).reshape(n, -1) for a in (input_ids, segment_ids, input_mask, positions, ids, nsl)]
    out = np.empty((n, sum(c.shape[1] for c in cols)), dtype=np.int32)
    at = 0
    for c in cols:
        out[:, at:at + c.shape[1]] = c
        at += c.shape[1]
    return out


def assemble_host(raw: torch.Tensor, L: int, P: int) -> List[torch.Tensor]:
    """CPU restatement of the device kernel (CPU runs and the test oracle)."""
    raw = raw.to(torch.int64)
    B = raw.shape[0]
    ids, seg, mask = raw[:, :L].clone(), raw[:, L:2 * L].clone(), raw[:, 2 * L:3 * L].clone()
    pos, mid = raw[:, 3 * L:3 * L + P], raw[:, 3 * L + P:3 * L + 2 * P]
    labels = torch.full((B, L), -1, dtype=torch.int64)
    live = torch.cumprod((pos != 0).to(torch.int64), dim=1).bool()
    for r in range(B):                                  # row loop keeps "last duplicate wins" (NumPy semantics)
        k = int(live[r].sum())
        labels[r, pos[r, :k]] = mid[r, :k]
    return [ids, seg, mask, labels, raw[:, 3 * L + 2 * P].clone()]


class DeviceBatchAssembler:
    """H2D of the packed matrix on a copy stream + the assembly kernel on the training stream."""

    def __init__(self, device: Optional[torch.device]):
        self.device = device
        self.cuda = device is not None and torch.device(device).type == "cuda"
        self._copy_stream = torch.cuda.Stream(device=device) if self.cuda else None

    def assemble(self, packed: PackedBertBatch) -> List[torch.Tensor]:
        L, P = packed.dims
        raw = packed.raw
        if not self.cuda:
            return assemble_host(raw, L, P)
        from ..ops import kernels
        cur = torch.cuda.current_stream(self.device)
        with torch.cuda.stream(self._copy_stream):
            raw_dev = raw.to(self.device, non_blocking=True)
        cur.wait_stream(self._copy_stream)
        raw_dev.record_stream(cur)
        return kernels.assemble_bert_batch(raw_dev, L, P)
