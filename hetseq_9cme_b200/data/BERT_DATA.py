"""In-memory BERT corpus (role of the reference's legacy hetseq/data/BERT_DATA.py:12-120 ``CombineBertData``, which
it abandoned because loading "7 % took 20 GB and ~25 min").  Kept for small corpora and tests: all shards are read
once into concatenated NumPy arrays; samples come back in the raw six-key order of the shard schema.  For training
prefer ``ConBertH5pyData`` (memory-mapped, vectorised batch fetch)."""
from __future__ import annotations

import collections

import numpy as np
import torch.utils.data

from . import shard_io


class CombineBertData(torch.utils.data.Dataset):
    def __init__(self, files, max_pred_length=512, keys=shard_io.KEYS):
        self.max_pred_length = max_pred_length
        self.keys = tuple(keys)
        cols = collections.OrderedDict((k, []) for k in self.keys)
        for path in files:
            sh = shard_io.open_shard(path)
            for k in self.keys:
                cols[k].append(np.asarray(sh[k]))
        self.inputs = collections.OrderedDict((k, np.concatenate(v)) for k, v in cols.items())

    def __len__(self):
        return len(self.inputs[self.keys[0]])

    def __getitem__(self, index):
        return [self.inputs[k][index] for k in self.keys]
