"""Learning-rate schedules (behavioural parity with hetseq/lr_scheduler.py:6-104).

Quirks kept on purpose: with ``--warmup-updates > 0`` the schedule is evaluated at
``num_updates = 0`` when the trainer is built, so the very first optimizer step runs with ``lr = 0``.
"""
from __future__ import annotations

from .optim import _Optimizer


class _LRScheduler:
    """Base: tracks the best validation loss; subclasses set the optimizer lr."""

    def __init__(self, args, optimizer):
        if not isinstance(optimizer, _Optimizer):
            raise ValueError("optimizer must be an instance of _Optimizer")
        self.args = args
        self.optimizer = optimizer
        self.best = None

    def state_dict(self):
        return {"best": self.best}

    def load_state_dict(self, state_dict):
        self.best = state_dict.get("best")

    def step(self, epoch, val_loss=None):
        """End-of-epoch hook."""
        if val_loss is not None:
            self.best = val_loss if self.best is None else min(self.best, val_loss)

    def step_update(self, num_updates):
        """After-every-update hook; returns the lr now in effect."""
        return self.optimizer.get_lr()


def polynomial_lr(num_updates: int, base_lr: float, warmup: int, total: int, end_lr: float, power: float) -> float:
    """Closed form of the schedule: linear warm-up to ``base_lr`` over ``warmup`` updates, then
    ``(base-end) * (1 - progress)**power + end`` until ``total`` updates, ``end_lr`` afterwards."""
    if warmup > 0 and num_updates <= warmup:
        return base_lr * (num_updates / float(warmup))
    if num_updates >= total:
        return end_lr
    remaining = 1.0 - (num_updates - warmup) / float(total - warmup)
    return (base_lr - end_lr) * remaining ** power + end_lr


class PolynomialDecayScheduler(_LRScheduler):
    """Polynomial decay with linear warm-up; per-epoch base lr from the ``--lr`` list."""

    def __init__(self, args, optimizer):
        super().__init__(args, optimizer)
        args.warmup_updates = getattr(args, "warmup_updates", 0) or 0
        self.lr = args.lr[0]
        self.warmup_factor = 1.0 / args.warmup_updates if args.warmup_updates > 0 else 1
        self.end_learning_rate = args.end_learning_rate
        self.total_num_update = args.total_num_update
        self.power = args.power
        self.optimizer.set_lr(self.warmup_factor * self.lr)

    def get_next_lr(self, epoch):
        lrs = self.args.lr
        force_anneal = getattr(self.args, "force_anneal", None)
        if force_anneal is None or epoch < force_anneal:
            return lrs[min(epoch, len(lrs) - 1)]
        return self.optimizer.get_lr()

    def step(self, epoch, val_loss=None):
        super().step(epoch, val_loss)
        self.lr = self.get_next_lr(epoch)
        self.optimizer.set_lr(self.warmup_factor * self.lr)
        return self.optimizer.get_lr()

    def step_update(self, num_updates):
        warmup = self.args.warmup_updates
        if warmup > 0 and num_updates <= warmup:
            self.warmup_factor = num_updates / float(warmup)
        lr = polynomial_lr(num_updates, self.lr, warmup, self.total_num_update, self.end_learning_rate, self.power)
        self.optimizer.set_lr(lr)
        return self.optimizer.get_lr()


# name (value of --lr-scheduler) -> class; extensions add theirs here and their flags to options._SCHED_ROWS
SCHEDULER_REGISTRY = {"PolynomialDecayScheduler": PolynomialDecayScheduler}


def build_lr_scheduler(args, optimizer):
    name = getattr(args, "lr_scheduler", "PolynomialDecayScheduler")
    if name not in SCHEDULER_REGISTRY:
        raise ValueError("unsupported lr_scheduler: {}".format(name))
    return SCHEDULER_REGISTRY[name](args, optimizer)
