"""Training meters (API parity with hetseq/meters.py:4-65: AverageMeter, TimeMeter, StopwatchMeter).

Wall-clock meters use ``time.perf_counter`` (monotonic) instead of ``time.time``; every meter can be
serialised with ``state_dict``/``load_state_dict`` so checkpoints can carry them as plain dicts.
"""
from __future__ import annotations

import time


class AverageMeter:
    """Running weighted mean: ``update(val, n)`` adds ``n`` observations of value ``val``."""

    __slots__ = ("val", "avg", "sum", "count")

    def __init__(self):
        self.reset()

    def reset(self):
        self.val = 0
        self.sum = 0
        self.count = 0
        self.avg = 0

    def update(self, val, n=1):
        self.val = val
        self.sum += val * n
        self.count += n
        self.avg = self.sum / self.count if self.count else 0

    def state_dict(self):
        return {"val": self.val, "sum": self.sum, "count": self.count}

    def load_state_dict(self, sd):
        self.val, self.sum, self.count = sd["val"], sd["sum"], sd["count"]
        self.avg = self.sum / self.count if self.count else 0


class TimeMeter:
    """Events per second since the last reset (``init`` seconds are credited up front)."""

    def __init__(self, init=0):
        self.reset(init)

    def reset(self, init=0):
        self.init = init
        self.start = time.perf_counter()
        self.n = 0

    def update(self, val=1):
        self.n += val

    @property
    def elapsed_time(self):
        return self.init + (time.perf_counter() - self.start)

    @property
    def avg(self):
        dt = self.elapsed_time
        return self.n / dt if dt > 0 else 0.0

    def state_dict(self):
        return {"init": self.elapsed_time, "n": self.n}

    def load_state_dict(self, sd):
        self.reset(init=sd.get("init", 0))
        self.n = sd.get("n", 0)


class StopwatchMeter:
    """Accumulated duration of start()/stop() intervals."""

    def __init__(self):
        self.reset()

    def reset(self):
        self.sum = 0
        self.n = 0
        self.start_time = None

    def start(self):
        self.start_time = time.perf_counter()

    def stop(self, n=1):
        if self.start_time is None:
            return
        self.sum += time.perf_counter() - self.start_time
        self.n += n
        self.start_time = None

    @property
    def avg(self):
        return self.sum / self.n if self.n else 0.0

    def state_dict(self):
        return {"sum": self.sum, "n": self.n}

    def load_state_dict(self, sd):
        self.sum, self.n, self.start_time = sd["sum"], sd["n"], None
