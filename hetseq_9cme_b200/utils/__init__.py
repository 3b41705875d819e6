"""Small helpers shared across the framework (live subset of hetseq/utils.py:12-37,86-91,167-171)."""
from __future__ import annotations

import math
from typing import Any, Callable

import torch


def apply_to_sample(fn: Callable[[torch.Tensor], Any], sample):
    """Map ``fn`` over every tensor in a nested dict / list / tuple structure."""
    if sample is None:
        return sample
    if torch.is_tensor(sample):
        return fn(sample)
    if isinstance(sample, dict):
        return {k: apply_to_sample(fn, v) for k, v in sample.items()}
    if isinstance(sample, (list, tuple)):
        return type(sample)(apply_to_sample(fn, v) for v in sample) if not hasattr(sample, "_fields") \
            else type(sample)(*(apply_to_sample(fn, v) for v in sample))
    return sample


def sample_is_empty(sample) -> bool:
    if sample is None:
        return True
    try:
        return len(sample) == 0
    except TypeError:
        return False


def move_to_cuda(sample, device=None, non_blocking: bool = True):
    """Nested ``.cuda()``; asynchronous when the source is pinned (ref: hetseq/utils.py:32-37 is synchronous)."""
    if sample_is_empty(sample):
        return {}
    return apply_to_sample(lambda t: t.cuda(device=device, non_blocking=non_blocking), sample)


def move_to_device(sample, device, non_blocking: bool = True):
    if sample_is_empty(sample):
        return {}
    return apply_to_sample(lambda t: t.to(device=device, non_blocking=non_blocking), sample)


def pin_sample(sample):
    return apply_to_sample(lambda t: t.pin_memory() if not t.is_pinned() else t, sample)


def item(x):
    """Tensor / numpy scalar -> python number."""
    if hasattr(x, "item"):
        return x.item()
    if hasattr(x, "__getitem__") and not isinstance(x, (str, bytes)):
        return x[0]
    return x


def get_perplexity(loss, round_to: int = 2):
    """``2 ** loss`` (the reference logs base-2 losses, hetseq/utils.py:167-171)."""
    if loss is None:
        return 0.0
    try:
        return round(math.pow(2, loss), round_to)
    except OverflowError:
        return float("inf")


def count_parameters(model, trainable_only: bool = False) -> int:
    return sum(p.numel() for p in model.parameters() if p.requires_grad or not trainable_only)
