"""Progress / log-line writers (API parity with hetseq/progress_bar.py:13-138).

``simple`` prints exactly the reference's line shape
``| epoch 001:      1 / 439 loss=0.044, ppl=1.03, ...`` so scripts that grep logs keep working.
``json`` (new) prints / appends one JSON object per logged step.
"""
from __future__ import annotations

import json
from collections import OrderedDict
from numbers import Number

from .meters import AverageMeter, StopwatchMeter, TimeMeter


def format_stat(stat):
    """Meter/number -> display string (formats follow hetseq/progress_bar.py:34-44)."""
    if isinstance(stat, Number):
        return "{:g}".format(stat)
    if isinstance(stat, AverageMeter):
        return "{:.3f}".format(stat.avg)
    if isinstance(stat, TimeMeter):
        return "{:g}".format(round(stat.avg))
    if isinstance(stat, StopwatchMeter):
        return "{:.4f}".format(stat.sum)
    return stat


def _raw_stat(stat):
    if isinstance(stat, (AverageMeter, TimeMeter)):
        return stat.avg
    if isinstance(stat, StopwatchMeter):
        return stat.sum
    return stat


class progress_bar:
    """Base class: wraps an iterable, remembers the epoch prefix, formats stats."""

    def __init__(self, iterable, epoch=None, prefix=None):
        self.iterable = iterable
        self.offset = getattr(iterable, "offset", 0)
        self.epoch = epoch
        parts = []
        if epoch is not None:
            parts.append("| epoch {:03d}".format(epoch))
        if prefix is not None:
            parts.append("| {}".format(prefix))
        self.prefix = " ".join(parts)

    def __len__(self):
        return len(self.iterable)

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        return False

    def __iter__(self):
        raise NotImplementedError

    def log(self, stats, tag="", step=None):
        raise NotImplementedError

    def print(self, stats, tag="", step=None):
        raise NotImplementedError

    @staticmethod
    def _format_stats(stats):
        return OrderedDict((k, str(format_stat(v))) for k, v in stats.items())

    @staticmethod
    def _str_commas(stats):
        return ", ".join("{}={}".format(k, v.strip()) for k, v in stats.items())

    @staticmethod
    def _str_pipes(stats):
        return " | ".join("{} {}".format(k, v.strip()) for k, v in stats.items())


class noop_progress_bar(progress_bar):
    """Iterates without logging."""

    def __iter__(self):
        yield from self.iterable

    def log(self, stats, tag="", step=None):
        pass

    def print(self, stats, tag="", step=None):
        pass


class simple_progress_bar(progress_bar):
    """One text line every ``log_interval`` steps."""

    def __init__(self, iterable, epoch=None, prefix=None, log_interval=1000):
        super().__init__(iterable, epoch, prefix)
        self.log_interval = log_interval
        self.stats = None

    def __iter__(self):
        size = len(self.iterable)
        for i, obj in enumerate(self.iterable, start=self.offset):
            yield obj
            due = self.log_interval is not None and i > 0 and i % self.log_interval == 0
            if due and self.stats is not None:
                print("{}:  {:5d} / {:d} {}".format(self.prefix, i, size, self._str_commas(self.stats)), flush=True)

    def log(self, stats, tag="", step=None):
        self.stats = self._format_stats(stats)

    def print(self, stats, tag="", step=None):
        print("{} | {}".format(self.prefix, self._str_pipes(self._format_stats(stats))), flush=True)


class json_progress_bar(progress_bar):
    """One JSON object per logged step, to stdout or appended to ``path``."""

    def __init__(self, iterable, epoch=None, prefix=None, log_interval=1000, path=None):
        super().__init__(iterable, epoch, prefix)
        self.log_interval = log_interval
        self.path = path
        self.stats = None

    def _emit(self, record):
        line = json.dumps(record, sort_keys=True)
        if self.path:
            with open(self.path, "a") as f:
                f.write(line + "\n")
        else:
            print(line, flush=True)

    def __iter__(self):
        size = len(self.iterable)
        for i, obj in enumerate(self.iterable, start=self.offset):
            yield obj
            due = self.log_interval is not None and i > 0 and i % self.log_interval == 0
            if due and self.stats is not None:
                rec = OrderedDict(epoch=self.epoch, update=i, size=size)
                rec.update(self.stats)
                self._emit(rec)

    def log(self, stats, tag="", step=None):
        self.stats = OrderedDict((k, _raw_stat(v)) for k, v in stats.items())

    def print(self, stats, tag="", step=None):
        rec = OrderedDict(epoch=self.epoch, tag=tag)
        rec.update((k, _raw_stat(v)) for k, v in stats.items())
        self._emit(rec)


class tee_progress_bar(progress_bar):
    """Drives a primary bar and mirrors every log()/print() into secondary sinks (e.g. JSONL)."""

    def __init__(self, primary, *sinks):
        super().__init__(primary.iterable, primary.epoch)
        self.primary = primary
        self.sinks = sinks
        self.prefix = primary.prefix
        self._step = 0

    def __iter__(self):
        for obj in self.primary:
            yield obj

    def log(self, stats, tag="", step=None):
        self.primary.log(stats, tag, step)
        for s in self.sinks:
            s.log(stats, tag, step)
            if s.stats is not None:
                rec = OrderedDict(epoch=s.epoch, update=step)
                rec.update(s.stats)
                s._emit(rec)

    def print(self, stats, tag="", step=None):
        self.primary.print(stats, tag, step)
        for s in self.sinks:
            s.print(stats, tag, step)


def build_progress_bar(args, iterator, epoch=None, prefix=None, default="simple", no_progress_bar="none"):
    """Pick the writer from ``--log-format`` / ``--no-progress-bar`` (ref: hetseq/progress_bar.py:13-31)."""
    fmt = getattr(args, "log_format", None)
    if fmt is None:
        fmt = no_progress_bar if getattr(args, "no_progress_bar", False) else default
    interval = getattr(args, "log_interval", 1)
    if fmt == "none":
        bar = noop_progress_bar(iterator, epoch, prefix)
    elif fmt == "simple":
        bar = simple_progress_bar(iterator, epoch, prefix, interval)
    elif fmt == "json":
        bar = json_progress_bar(iterator, epoch, prefix, interval)
    else:
        raise ValueError("Unknown log format: {}".format(fmt))
    jsonl = getattr(args, "log_jsonl", None)
    if jsonl and (getattr(args, "distributed_rank", 0) or 0) == 0:      # one writer: the master rank, like the text log
        bar = tee_progress_bar(bar, json_progress_bar(iterator, epoch, prefix, interval, path=jsonl))
    return bar
