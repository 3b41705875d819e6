"""Task base class, BERT pre-training task and the MNIST example task
(API of hetseq/tasks/tasks.py:22-343; extension contract in the reference's docs/source/extending.rst:5-47).

A task owns the datasets, builds the model and defines what one forward/backward (``train_step``) and one
forward (``valid_step``) mean, including the ``sample_size`` that becomes the gradient denominator.
"""
from __future__ import annotations

import collections
import os
from typing import Dict, Optional

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from ..data import BertH5pyData, ConBertH5pyData, MNISTDataset, data_utils, iterators


class Task:
    def __init__(self, args):
        self.args = args
        self.datasets: Dict[str, torch.utils.data.Dataset] = {}
        self.dataset_to_epoch_iter = {}

    @classmethod
    def setup_task(cls, args, **kwargs):
        return cls(args)

    @staticmethod
    def load_dictionary(vocab_file):
        """``vocab.txt`` (one token per line) -> OrderedDict token -> index."""
        vocab = collections.OrderedDict()
        with open(vocab_file, "r", encoding="utf-8") as reader:
            for index, line in enumerate(reader):
                vocab[line.strip()] = index
        print("| loaded dictionary with {} subwords  from: {}".format(len(vocab), vocab_file))
        return vocab

    def load_dataset(self, split, **kwargs):
        raise NotImplementedError

    def dataset(self, split):
        if split not in self.datasets:
            raise KeyError("Dataset not loaded: " + split)
        if not isinstance(self.datasets[split], torch.utils.data.Dataset):
            raise TypeError("Datasets are expected to be of type torch.utils.data.Dataset")
        return self.datasets[split]

    def get_batch_iterator(self, dataset, max_tokens=None, max_sentences=None, max_positions=None,
                           ignore_invalid_inputs=False, required_batch_size_multiple=1, seed=1, num_shards=1, shard_id=0,
                           num_workers=0, epoch=0):
        """A reusable, sharded, resumable iterator over ``dataset`` (one instance per dataset object is cached,
        like the reference does, so epochs continue on the same iterator)."""
        if dataset in self.dataset_to_epoch_iter:
            return self.dataset_to_epoch_iter[dataset]
        with data_utils.numpy_seed(seed):
            indices = dataset.ordered_indices()
        print("| build batch sampler")
        sizes = dataset.sizes_array() if hasattr(dataset, "sizes_array") else None
        batch_sampler = data_utils.batch_by_size(indices, None if sizes is not None else dataset.num_tokens,
                                                 max_tokens=max_tokens, max_sentences=max_sentences,
                                                 required_batch_size_multiple=required_batch_size_multiple, sizes=sizes)
        print("| finish building batch sampler")
        use_cuda = torch.cuda.is_available() and not getattr(self.args, "cpu", False)
        epoch_iter = iterators.EpochBatchIterator(dataset=dataset, collate_fn=dataset.collater, batch_sampler=batch_sampler,
                                                  seed=seed, num_shards=num_shards, shard_id=shard_id,
                                                  num_workers=num_workers, epoch=epoch, pin_memory=use_cuda)
        self.dataset_to_epoch_iter[dataset] = epoch_iter
        return epoch_iter

    def build_model(self, args):
        raise NotImplementedError

    # ---- what a step means ---------------------------------------------------------------------------
    def _forward(self, sample, model):
        return model(*sample)

    def _sample_size(self, sample) -> int:
        """The reference's denominator: ``len(sample[0][0])`` — sequence length for BERT, 1 for MNIST
        (hetseq/tasks/tasks.py:170-173)."""
        if sample is None or len(sample) == 0:
            return 0
        first = sample[0]
        if len(first) == 0 or len(first[0]) == 0:
            return 0
        return len(first[0])

    def train_step(self, sample, model, optimizer, ignore_grad=False):
        model.train()
        loss = self._forward(sample, model)
        if ignore_grad:
            loss = loss * 0
        sample_size = self._sample_size(sample)
        logging_output = {"nsentences": sample_size, "loss": loss.detach(), "nll_loss": loss.detach(),
                          "ntokens": self._num_tokens(sample), "sample_size": sample_size}
        optimizer.backward(loss)
        return loss, sample_size, logging_output

    @torch.no_grad()
    def valid_step(self, sample, model):
        model.eval()
        loss = self._forward(sample, model)
        sample_size = self._sample_size(sample)
        return loss, sample_size, {"nsentences": sample_size, "loss": loss.detach(), "nll_loss": loss.detach(),
                                   "ntokens": self._num_tokens(sample), "sample_size": sample_size}

    def _num_tokens(self, sample) -> int:
        return 0

    def update_step(self, num_updates):
        """Hook called after every optimizer step."""


class LanguageModelingTask(Task):
    """BERT pre-training (phase 1: seq 128, phase 2: seq 512) on pre-tokenised shards."""

    def __init__(self, args, dictionary):
        super().__init__(args)
        self.dictionary = dictionary

    @classmethod
    def setup_task(cls, args, **kwargs):
        dictionary = cls.load_dictionary(args.dict) if getattr(args, "dict", None) else collections.OrderedDict()
        return cls(args, dictionary)

    def build_model(self, args):
        if args.task != "bert":
            raise ValueError("Unsupported language modeling task: {}".format(args.task))
        from ..bert_modeling import BertConfig, BertForPreTraining
        config = BertConfig.from_json_file(args.config_file)
        model = BertForPreTraining(config)
        # capacity of the labelled-positions MLM head: never below the width of the shards' masked_lm_positions (a
        # smaller flag would silently drop labels — e.g. "20" from a phase-1 script on phase-2 shards carrying 80)
        flag, width = int(getattr(args, "max_predictions_per_seq", 0) or 0), int(getattr(self, "_max_pred", 0) or 0)
        if flag and width and flag < width:
            print("| WARNING: --max-predictions-per-seq {} is below the data's masked_lm_positions width {}; using {}".format(
                flag, width, width))
        model.max_predictions_per_seq = max(flag, width)
        model.checkpoint_activations = bool(getattr(args, "checkpoint_activations", False))
        return model

    def load_dataset(self, split, **kwargs):
        path = self.args.data
        if not os.path.exists(path):
            raise FileNotFoundError("Dataset not found: ({})".format(path))
        files = [os.path.join(path, f) for f in os.listdir(path)] if os.path.isdir(path) else [path]
        files = sorted(f for f in files if split in os.path.basename(f) and not f.endswith(".tmp"))
        if getattr(self.args, "num_file", 0) > 0:
            files = files[: self.args.num_file]
        assert len(files) > 0, "no suitable file in split ***{}***".format(split)
        shards = [BertH5pyData(f, getattr(self.args, "max_pred_length", 512)) for f in files]
        dataset = ConBertH5pyData(shards)
        dataset.packed = bool(getattr(self.args, "device_batch_assembly", False))
        # width of masked_lm_positions bounds the labelled positions per sequence -> capacity of the MLM head
        try:
            self._max_pred = int(shards[0]._shard["masked_lm_positions"].shape[1])
        except Exception:
            self._max_pred = 0
        print("| loaded {} sentences from: {}".format(len(dataset), path), flush=True)
        self.datasets[split] = dataset
        print("| loading finished")

    def _num_tokens(self, sample) -> int:
        try:
            return int(sample[0].numel())
        except Exception:
            return 0


class MNISTNet(nn.Module):
    """conv(1,32,3) - relu - conv(32,64,3) - relu - maxpool2 - drop2d(.25) - fc(9216,128) - relu - drop(.5) -
    fc(128,10) - log_softmax - nll; ``forward(x, target)`` returns the loss (hetseq/tasks/tasks.py:318-343)."""

    def __init__(self):
        super().__init__()
        self.conv1 = nn.Conv2d(1, 32, 3, 1)
        self.conv2 = nn.Conv2d(32, 64, 3, 1)
        no_drop = bool(os.environ.get("B200SEQ_TEST_NO_DROPOUT"))      # test hook: deterministic replicas
        self.dropout1 = nn.Dropout2d(0.0 if no_drop else 0.25)
        self.dropout2 = nn.Dropout(0.0 if no_drop else 0.5)
        self.fc1 = nn.Linear(9216, 128)
        self.fc2 = nn.Linear(128, 10)

    def logits(self, x):
        x = F.relu(self.conv1(x))
        x = F.max_pool2d(F.relu(self.conv2(x)), 2)
        x = torch.flatten(self.dropout1(x), 1)
        x = self.dropout2(F.relu(self.fc1(x)))
        return self.fc2(x)

    def forward(self, x, target=None, eval=False):
        out = F.log_softmax(self.logits(x), dim=1)
        if target is None or eval:
            return out
        return F.nll_loss(out, target)


class MNISTTask(Task):
    def build_model(self, args):
        return MNISTNet()

    def load_dataset(self, split, **kwargs):
        path = self.args.data
        if not os.path.exists(path):
            raise FileNotFoundError("Dataset not found: ({})".format(path))
        if os.path.isdir(path):
            proc = os.path.join(path, "MNIST", "processed")
            if os.path.exists(proc):
                path = proc
            elif os.path.basename(os.path.normpath(path)) != "processed":
                raise FileNotFoundError("{} has no MNIST/processed/ directory (no network here to download it; "
                                        "see data/synthetic.make_mnist)".format(path))
        files = [os.path.join(path, f) for f in os.listdir(path)] if os.path.isdir(path) else [path]
        files = sorted(f for f in files if split in os.path.basename(f))
        assert len(files) == 1, "no suitable file in split ***{}***".format(split)
        dataset = MNISTDataset(files[0])
        print("| loaded {} sentences from: {}".format(len(dataset), path), flush=True)
        self.datasets[split] = dataset
        print("| loading finished")

    def _num_tokens(self, sample) -> int:
        try:
            return int(sample[0].size(0))
        except Exception:
            return 0
