"""Data-parallel gradient reduction over the flat gradient arena.

The reference delegates this to ``torch.nn.parallel.DistributedDataParallel`` (hetseq/controller.py:75-90): its
Reducer copies gradients into 25 MB buckets, all-reduces each with NCCL, divides by the world size and copies
back.  Here gradients already live in one flat fp32 buffer, buckets are slices of it, and a reducer is a
``GradSink``: backward code tells it which parameters are final and it launches the reduction of a bucket as
soon as all of the bucket's parameters are in, on a side stream, while backward keeps running.

Two back ends:

* ``TorchReducer`` — ``torch.distributed.all_reduce(SUM)`` per bucket (NCCL on GPUs, gloo on CPU).  The baseline
  and the portable path (multi-node, CPU tests).
* ``P2PReducer`` (``parallel/p2p.py``) — hand-written sm_100a kernels that read and write the peers' arenas
  directly over NVLink (two-shot reduce-scatter + all-gather, one-shot for small buckets).

Both produce the SUM over ranks; the 1/sample_size scale and the clip coefficient are applied inside the fused
optimizer kernel, so no separate "divide by world size" pass exists.
"""
from __future__ import annotations

from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..models.bert_native import GradSink


class BaseReducer(GradSink):
    def __init__(self, arena, world_size: int):
        self.arena = arena
        self.world_size = world_size
        self._pending: List[int] = []
        self._remaining: Dict[int, int] = {}
        self._active = False
        self._expected = {bi: len({id(p) for p in ms}) for bi, (_, _, ms) in enumerate(arena.buckets)}
        self._seen = set()
        self._after_backward = False
        self.fused_norm = False           # True: reductions leave sum(g^2) in arena.tail (no separate norm pass)
        self.sharded = False              # True: reduce-scatter only; the optimizer step runs on owned slices (p2p.py)

    # ---- step protocol ---------------------------------------------------------------------------
    def begin(self, sync_now: bool):
        """Called before each micro-batch's backward; ``sync_now`` is True only for the last one."""
        self._active = sync_now and self.world_size > 1
        self._remaining = dict(self._expected)
        self._seen = set()
        self._launched = set()

    def ready(self, params) -> None:
        if not self._active:
            return
        for p in params:
            pid = id(p)
            if pid in self._seen or pid not in self.arena.bucket_of:
                continue
            self._seen.add(pid)
            bi = self.arena.bucket_of[pid]
            self._remaining[bi] -= 1
            if self._remaining[bi] == 0:
                self._launch(bi)

    def enable_timing(self, on: bool = True):
        """Measure, per step, how long the training stream has to wait for reductions after backward is done
        (the *exposed* all-reduce time).  Uses timing events, so it cannot be on while a CUDA graph is captured."""
        self._timing = on and self.arena.device.type == "cuda"
        self._timing_pairs = []

    def exposed_ms(self):
        """Mean exposed wait per timed step (call after a synchronize)."""
        pairs = getattr(self, "_timing_pairs", [])
        if not pairs:
            return None
        vals = [max(0.0, a.elapsed_time(b)) for a, b in pairs]
        return sum(vals) / len(vals)

    def finish(self):
        """After backward: reduce whatever was not announced (unused parameters, torch-autograd-only models),
        then make the training stream wait for all reductions."""
        timing = getattr(self, "_timing", False) and self._active
        if timing:
            e_bwd = torch.cuda.Event(enable_timing=True)
            e_bwd.record(torch.cuda.current_stream())          # backward is complete at this point of the stream
        if self._active:
            self._after_backward = True           # nothing competes for the SMs any more: back ends may go wide
            for bi in range(len(self.arena.buckets)):
                if bi not in self._launched:
                    self._launch(bi)
            self._after_backward = False
            self._wait_all()
            if timing:
                e_red = torch.cuda.Event(enable_timing=True)
                e_red.record(torch.cuda.current_stream())      # ... and here every reduction has landed
                self._timing_pairs.append((e_bwd, e_red))
        self._active = False

    # ---- back-end hooks --------------------------------------------------------------------------
    def _launch(self, bucket_index: int):
        raise NotImplementedError

    def _wait_all(self):
        raise NotImplementedError

    def reduce_tail(self):
        """Sum the statistics vector (``arena.tail``) over ranks, in stream order on the training stream."""
        if self.world_size > 1:
            dist.all_reduce(self.arena.tail)

    def gather_sharded(self, flats) -> None:
        """Sharded back ends make tensors whose slices only their owners keep current whole again (checkpoints)."""

    def broadcast_parameters(self):
        """Make every rank start from rank 0's parameters (what DDP does at construction)."""
        if self.world_size > 1:
            dist.broadcast(self.arena.master, src=0)
            if self.arena.work is not None:
                self.arena.sync_work_from_master()


class TorchReducer(BaseReducer):
    def __init__(self, arena, world_size: int, group=None):
        super().__init__(arena, world_size)
        self.group = group
        self._works = []
        self._cuda = arena.device.type == "cuda"
        self._side = torch.cuda.Stream(device=arena.device) if self._cuda else None

    def _launch(self, bi: int):
        self._launched.add(bi)
        s, e, _ = self.arena.buckets[bi]
        view = self.arena.grad[s:e]
        if self._cuda:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream())
            with torch.cuda.stream(self._side):
                self._side.wait_event(ev)
                self._works.append(dist.all_reduce(view, group=self.group, async_op=True))
        else:
            self._works.append(dist.all_reduce(view, group=self.group, async_op=True))

    def _wait_all(self):
        for w in self._works:
            w.wait()
        self._works = []
        if self._cuda:
            torch.cuda.current_stream().wait_stream(self._side)


class NullReducer(BaseReducer):
    """world_size == 1."""

    def _launch(self, bi):
        self._launched.add(bi)

    def _wait_all(self):
        pass


class LocalNormReducer(BaseReducer):
    """world_size == 1 on a GPU: nothing to reduce, but the clip needs sum(g^2) over 440 MB of gradients — a 67 us
    memory-bound pass at the end of backward.  Here each bucket's share is summed on a side stream as soon as backward
    has produced it (same ``ready`` notifications the multi-GPU reducers use), next to the tensor-core kernels that are
    still running, and lands in ``arena.tail[SQ_SLOT]`` like the reducers' fused norm.  Measured on B200
    (profiles/README.md §18): 18 small launches with their fork / join edges in the captured graph are slower than the
    one pass they replace (7.89 vs 7.70 ms / step) — opt-in with B200_LOCAL_NORM=1."""

    def __init__(self, arena):
        super().__init__(arena, 1)
        self.fused_norm = True
        self._side = torch.cuda.Stream(device=arena.device)
        from .p2p import SQ_SLOT
        self._sq = arena.tail[SQ_SLOT:SQ_SLOT + 1]

    def begin(self, sync_now: bool):
        super().begin(sync_now)
        self._active = sync_now            # the base class only goes active for world_size > 1

    def _launch(self, bi):
        from ..ops import kernels
        self._launched.add(bi)
        s, e, _ = self.arena.buckets[bi]
        # always on the side stream: the launches share one scratch buffer and must not overlap each other
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self._side.wait_event(ev)
        with torch.cuda.stream(self._side):
            kernels.sumsq(self.arena.grad[s:e], self._sq, accumulate=True)

    def _wait_all(self):
        torch.cuda.current_stream().wait_stream(self._side)

    def reduce_tail(self):
        pass


def build_reducer(args, arena, world_size: int, sharded: bool = False):
    """Pick the back end.  Every rank must end up with the SAME kind of reducer (they issue different collectives),
    so a failure on one rank is agreed on by all before anyone falls back."""
    if world_size <= 1:
        import os
        if arena.device.type == "cuda" and os.environ.get("B200_LOCAL_NORM"):
            return LocalNormReducer(arena)
        return NullReducer(arena, 1)
    comm = getattr(args, "comm", "auto")
    if arena.device.type == "cuda" and comm in ("auto", "p2p"):
        from . import p2p
        red, err = None, None
        try:
            red = p2p.P2PReducer(arena, world_size, sharded=sharded)
        except Exception as e:  # peer access unavailable (multi-node, MIG, ...) -> documented fallback
            err = e
        if p2p._all_ranks_ok(red is not None):
            return red
        if red is not None:
            red.close()
        if comm == "p2p":
            raise RuntimeError("--comm p2p: the peer-to-peer reducer could not be set up on every rank ({})".format(err or "peer failure"))
        from ..distributed_utils import log_all
        log_all("| p2p reducer unavailable ({}); using torch.distributed all-reduce".format(err or "a peer rank failed"))
    return TorchReducer(arena, world_size)
