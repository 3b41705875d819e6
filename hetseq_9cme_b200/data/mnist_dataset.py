"""MNIST from torchvision's processed ``.pt`` tuples (API of hetseq/data/mnist_dataset.py:11-76).

Normalisation ``(x/255 - 0.1307) / 0.3081`` is applied with tensor arithmetic on whole batches
(``fetch_batch``) instead of a PIL round trip per image."""
from __future__ import annotations

import numpy as np
import torch
import torch.utils.data

MEAN, STD = 0.1307, 0.3081


class MNISTDataset(torch.utils.data.Dataset):
    def __init__(self, path):
        super().__init__()
        self.path = path
        self.images, self.labels = torch.load(path, weights_only=False)
        self.images = torch.as_tensor(self.images)
        self.labels = torch.as_tensor(self.labels).long()
        self._len = len(self.labels)

    def __len__(self):
        return self._len

    def _norm(self, img_u8):
        return (img_u8.float().div_(255.0) - MEAN) / STD

    def __getitem__(self, index):
        if index < 0 or index >= self._len:
            raise IndexError("index out of range")
        return self._norm(self.images[index]).unsqueeze(0), self.labels[index]

    def fetch_batch(self, indices):
        idx = torch.as_tensor(np.asarray(indices, dtype=np.int64))
        return [self._norm(self.images[idx]).unsqueeze(1), self.labels[idx]]

    def collater(self, samples):
        if len(samples) == 0:
            return None
        return torch.utils.data.default_collate(samples)

    def ordered_indices(self):
        return np.arange(len(self))

    def num_tokens(self, index):
        return 1

    def sizes_array(self):
        return np.ones(len(self), dtype=np.int64)

    def size(self, index):
        return 1

    def set_epoch(self, epoch):
        pass
