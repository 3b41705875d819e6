"""Token-classification dataset wrappers (API of hetseq/data/bert_ner_dataset.py:7-34 and bert_el_dataset.py).

Wrap any indexable collection of feature dicts (an HF ``datasets.Dataset`` or a plain list); batches are
padded to the longest sequence in the batch by the collator in ``args.data_collator``."""
from __future__ import annotations

import numpy as np
import torch.utils.data


class BertNerDataset(torch.utils.data.Dataset):
    def __init__(self, dataset, args):
        super().__init__()
        self.dataset = dataset
        self.args = args
        self.collate_fn = getattr(args, "data_collator", None)

    def __len__(self):
        return len(self.dataset)

    def __getitem__(self, index):
        return self.dataset[int(index)]

    def ordered_indices(self):
        return np.arange(len(self.dataset))

    def num_tokens(self, index):
        return len(self.dataset[int(index)]["labels"])

    def sizes_array(self):
        return np.fromiter((len(self.dataset[i]["labels"]) for i in range(len(self.dataset))), dtype=np.int64,
                           count=len(self.dataset))

    def collater(self, samples):
        if len(samples) == 0:
            return None
        return self.collate_fn(samples)

    def set_epoch(self, epoch):
        pass


class BertELDataset(BertNerDataset):
    """Same contract with an extra ``entity_labels`` feature (entity-linking task)."""
