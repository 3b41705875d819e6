"""Epoch / shard / group iterators (semantics of hetseq/data/iterators.py:10-275).

Kept from the reference, because resumable training and cross-rank lock-step depend on them:

* batch *composition* is frozen once; each epoch only the batch *order* is reshuffled with NumPy's legacy
  generator seeded with ``seed + epoch`` (so a resumed run sees the same order);
* rank ``r`` of ``W`` takes batches ``r, r+W, ...`` and short ranks are padded with ``[]`` so every rank runs
  the same number of steps;
* ``state_dict`` = ``{'epoch', 'iterations_in_epoch'}`` and loading it fast-forwards inside the epoch.

Different from the reference: batches are produced by a background *thread* that assembles whole batches
(``dataset.fetch_batch`` when the dataset offers it), optionally pins them, and runs ahead of the training
loop by ``prefetch`` batches — no worker processes, no per-sample file opens.
"""
from __future__ import annotations

import itertools
import math
import queue
import threading
from typing import Callable, List, Optional, Sequence

import numpy as np
import torch

from . import data_utils


class CountingIterator:
    """Wraps an iterable and counts consumed elements (``count``), supports ``skip`` and ``has_next``."""

    def __init__(self, iterable, start: int = 0, length: Optional[int] = None):
        self.iterable = iterable
        self.count = start
        self.len = start + (len(iterable) if length is None else length)
        self.itr = iter(self)

    def __len__(self):
        return self.len

    def __iter__(self):
        for x in self.iterable:
            self.count += 1
            yield x

    def __next__(self):
        return next(self.itr)

    def has_next(self) -> bool:
        return self.count < len(self)

    def skip(self, num_to_skip: int):
        next(itertools.islice(self.itr, num_to_skip, num_to_skip), None)
        return self


class ShardedIterator:
    """Every ``num_shards``-th element starting at ``shard_id``, padded with ``fill_value`` to equal length."""

    def __init__(self, iterable, num_shards: int, shard_id: int, fill_value=None):
        if shard_id < 0 or shard_id >= num_shards:
            raise ValueError("shard_id must be between 0 and num_shards")
        n = len(iterable)
        self._sharded_len = (n + num_shards - 1) // num_shards
        own = itertools.islice(iterable, shard_id, n, num_shards)
        self.itr = itertools.zip_longest(range(self._sharded_len), own, fillvalue=fill_value)

    def __len__(self):
        return self._sharded_len

    def __iter__(self):
        return self

    def __next__(self):
        return next(self.itr)[1]


class GroupedIterator:
    """Chunks of ``chunk_size`` consecutive items (the micro-batches of one optimizer step)."""

    def __init__(self, iterable, chunk_size: int):
        self._len = int(math.ceil(len(iterable) / float(chunk_size)))
        self.offset = int(math.ceil(getattr(iterable, "count", 0) / float(chunk_size)))
        self.itr = iterable
        self.chunk_size = chunk_size

    def __len__(self):
        return self._len

    def __iter__(self):
        return self

    def __next__(self):
        chunk = []
        try:
            for _ in range(self.chunk_size):
                chunk.append(next(self.itr))
        except StopIteration:
            if not chunk:
                raise
        return chunk


class BatchLoader:
    """Materialises batches (lists of indices) into collated samples, ``prefetch`` batches ahead on a thread."""

    _END = object()

    def __init__(self, dataset, collate_fn: Callable, batches: Sequence[Sequence[int]], prefetch: int = 2,
                 pin_memory: bool = False):
        self.dataset = dataset
        self.collate_fn = collate_fn
        self.batches = list(batches)
        self.prefetch = max(0, int(prefetch))
        self.pin_memory = pin_memory and torch.cuda.is_available()

    def __len__(self):
        return len(self.batches)

    def _make(self, idxs):
        if len(idxs) == 0:
            return self.collate_fn([])
        fetch = getattr(self.dataset, "fetch_batch", None)
        sample = fetch(idxs) if fetch is not None else self.collate_fn([self.dataset[i] for i in idxs])
        if self.pin_memory and sample is not None:
            sample = self._pin(sample)
        return sample

    # Pinned staging ring: batches are copied into page-locked buffers that are allocated once per tensor signature
    # instead of calling Tensor.pin_memory() per batch.  That call goes to the caching host allocator from this thread
    # and now and then ends in cudaHostAlloc while holding the GIL — measured on B200 as a 20 ms bubble in one training
    # step out of ~40 (profiles/README.md §8).  A buffer is reused only when (a) RING_EXTRA + prefetch newer batches
    # have been produced since — the controller consumes statistics in lock step, at most one step behind the GPU, so
    # the H2D copy that read it has long completed — and (b) no consumer still holds the tensor (reference count), which
    # covers --update-freq groups and callers that keep batches around; otherwise the ring grows.
    RING_EXTRA = 6
    RING_MAX_BUFFERS = 256

    def _pin(self, sample):
        import collections
        import sys
        from ..utils import apply_to_sample, pin_sample
        if not hasattr(self, "_ring"):
            self._ring, self._ring_total = {}, 0
        depth = self.prefetch + self.RING_EXTRA

        def stage(t):
            if not torch.is_tensor(t) or t.is_pinned() or t.numel() == 0:
                return t
            key = (tuple(t.shape), t.dtype)
            q = self._ring.setdefault(key, collections.deque())
            buf = None
            for j in range(len(q) - depth + 1):                      # old enough: >= depth - 1 newer batches exist
                if sys.getrefcount(q[j]) == 2:                       # held by the deque only (+ this call's argument)
                    buf = q[j]
                    del q[j]
                    break
            if buf is not None:
                pass
            elif self._ring_total < self.RING_MAX_BUFFERS:
                buf = torch.empty(t.shape, dtype=t.dtype).pin_memory()
                self._ring_total += 1
            else:                                                    # ragged shapes / hoarding consumer: allocator path
                return t.pin_memory()
            buf.copy_(t)
            q.append(buf)
            return buf

        try:
            return apply_to_sample(stage, sample)
        except RuntimeError:
            return pin_sample(sample)

    def __iter__(self):
        if self.prefetch == 0:
            for b in self.batches:
                yield self._make(b)
            return
        q: "queue.Queue" = queue.Queue(maxsize=self.prefetch)
        stop = threading.Event()

        def work():
            try:
                for b in self.batches:
                    if stop.is_set():
                        return
                    q.put(self._make(b))
                q.put(self._END)
            except BaseException as e:  # surface loader errors in the consumer
                q.put(e)

        t = threading.Thread(target=work, daemon=True, name="b200seq-loader")
        t.start()
        try:
            while True:
                item = q.get()
                if item is self._END:
                    return
                if isinstance(item, BaseException):
                    raise item
                yield item
        finally:
            stop.set()
            while not q.empty():
                try:
                    q.get_nowait()
                except queue.Empty:
                    break


class EpochBatchIterating:
    def __len__(self) -> int:
        raise NotImplementedError

    def next_epoch_itr(self, shuffle=True, fix_batches_to_gpus=False):
        raise NotImplementedError

    def end_of_epoch(self) -> bool:
        raise NotImplementedError

    @property
    def iterations_in_epoch(self) -> int:
        raise NotImplementedError

    def state_dict(self):
        raise NotImplementedError

    def load_state_dict(self, state_dict):
        raise NotImplementedError


class EpochBatchIterator(EpochBatchIterating):
    """Multi-epoch, shardable, resumable iterator over a dataset with a frozen batch sampler."""

    def __init__(self, dataset, collate_fn, batch_sampler, seed=1, num_shards=1, shard_id=0, num_workers=0, epoch=0,
                 pin_memory=False):
        self.dataset = dataset
        self.collate_fn = collate_fn
        self.frozen_batches = tuple(list(b) for b in batch_sampler)
        self.seed = seed
        self.num_shards = num_shards
        self.shard_id = shard_id
        self.num_workers = num_workers
        self.pin_memory = pin_memory
        self.epoch = epoch
        self._cur_epoch_itr = None
        self._next_epoch_itr = None
        self._shuffle = True

    def __len__(self):
        return len(self.frozen_batches)

    def next_epoch_itr(self, shuffle=True, fix_batches_to_gpus=False):
        if self._next_epoch_itr is not None:
            self._cur_epoch_itr, self._next_epoch_itr = self._next_epoch_itr, None
        else:
            self.epoch += 1
            self._cur_epoch_itr = self._get_iterator_for_epoch(self.epoch, shuffle, fix_batches_to_gpus)
        self._shuffle = shuffle
        if hasattr(self.dataset, "set_epoch"):
            self.dataset.set_epoch(self.epoch)
        return self._cur_epoch_itr

    def end_of_epoch(self) -> bool:
        return not self._cur_epoch_itr.has_next()

    @property
    def iterations_in_epoch(self) -> int:
        if self._cur_epoch_itr is not None:
            return self._cur_epoch_itr.count
        if self._next_epoch_itr is not None:
            return self._next_epoch_itr.count
        return 0

    def state_dict(self):
        return {"epoch": self.epoch, "iterations_in_epoch": self.iterations_in_epoch, "shuffle": self._shuffle}

    def load_state_dict(self, state_dict):
        self.epoch = state_dict["epoch"]
        pos = state_dict.get("iterations_in_epoch", 0)
        if pos > 0:
            self._next_epoch_itr = self._get_iterator_for_epoch(self.epoch, state_dict.get("shuffle", True), offset=pos)

    def batches_for_epoch(self, epoch, shuffle=True, fix_batches_to_gpus=False) -> List[List[int]]:
        """This rank's batch list for ``epoch`` (exposed for tests and for the lock-step logic)."""
        batches = list(self.frozen_batches)
        if shuffle and not fix_batches_to_gpus:
            with data_utils.numpy_seed(self.seed + epoch):
                np.random.shuffle(batches)
        mine = list(ShardedIterator(batches, self.num_shards, self.shard_id, fill_value=[]))
        if shuffle and fix_batches_to_gpus:
            with data_utils.numpy_seed(self.seed + epoch + self.shard_id):
                np.random.shuffle(mine)
        return mine

    def _get_iterator_for_epoch(self, epoch, shuffle, fix_batches_to_gpus=False, offset=0):
        mine = self.batches_for_epoch(epoch, shuffle, fix_batches_to_gpus)
        if offset > 0 and offset >= len(mine):
            return None
        loader = BatchLoader(self.dataset, self.collate_fn, mine[offset:], prefetch=max(2, self.num_workers) if self.num_workers != 0 else 0,
                             pin_memory=self.pin_memory)
        return CountingIterator(loader, start=offset)
