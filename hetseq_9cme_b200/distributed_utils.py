"""Process-group bootstrap and small collectives (API of hetseq/distributed_utils.py:11-132).

``torch.distributed`` is the control plane only (rendezvous, object exchange, the CPU/gloo path and the NCCL
baseline); the gradient hot path lives in ``parallel/reducer.py``.  Unlike the reference, nothing here assumes
CUDA: ``all_gather_list`` and the warm-up collective work on gloo/CPU, which is what makes the 2-process CPU
configuration (BASELINE config #1) runnable at all.
"""
from __future__ import annotations

import builtins
import os
import pickle
import socket
import warnings
from typing import Any, List, Optional

import torch
import torch.distributed as dist

_builtin_print = builtins.print


def is_master(args) -> bool:
    return getattr(args, "distributed_rank", 0) in (0, None)


def suppress_output(is_master: bool) -> None:
    """Silence ``print`` on non-master ranks; ``print(..., force=True)`` always prints."""

    def _print(*a, **kw):
        force = kw.pop("force", False)
        if is_master or force:
            _builtin_print(*a, **kw)

    builtins.print = _print


def log_all(*a, **kw) -> None:
    """Print on every rank regardless of ``suppress_output``."""
    kw.setdefault("flush", True)
    _builtin_print(*a, **kw)


def restore_output() -> None:
    builtins.print = _builtin_print


def _use_cuda(args) -> bool:
    return torch.cuda.is_available() and not getattr(args, "cpu", False)


def distributed_init(args) -> int:
    """Join the process group described by ``--distributed-*`` and return this process's rank."""
    if args.distributed_world_size == 1:
        raise ValueError("Cannot initialize distributed with distributed_world_size=1")
    if dist.is_initialized():
        warnings.warn("Distributed is already initialized, cannot initialize twice!")
    else:
        backend = args.distributed_backend
        if not _use_cuda(args) and backend == "nccl":
            backend = "gloo"
            args.distributed_backend = backend
        _builtin_print("| distributed init (rank {}): {}".format(args.distributed_rank, args.distributed_init_method), flush=True)
        dist.init_process_group(backend=backend, init_method=args.distributed_init_method,
                                world_size=args.distributed_world_size, rank=args.distributed_rank)
        _builtin_print("| initialized host {} as rank {}".format(socket.gethostname(), args.distributed_rank), flush=True)
        # one tiny collective creates the communicator before the first timed step
        warm = torch.zeros(1, device="cuda" if _use_cuda(args) else "cpu")
        dist.all_reduce(warm)
        suppress_output(is_master(args))
    args.distributed_rank = dist.get_rank()
    return args.distributed_rank


def get_rank() -> int:
    return dist.get_rank() if dist.is_available() and dist.is_initialized() else 0


def get_world_size() -> int:
    return dist.get_world_size() if dist.is_available() and dist.is_initialized() else 1


def get_default_group():
    return dist.group.WORLD


def all_reduce(tensor, group=None):
    if group is None:
        group = get_default_group()
    return dist.all_reduce(tensor, group=group)


def barrier():
    if dist.is_available() and dist.is_initialized():
        dist.barrier()


def all_gather_list(data: Any, group=None, max_size: int = 16384) -> List[Any]:
    """Gather one picklable object per rank.

    Same contract as the reference (fixed ``max_size`` slot per rank, every rank must call it), implemented as a
    real all-gather of length-prefixed byte slots on the backend's device instead of the reference's
    sum-all-reduce over a zeroed ``max_size * world`` CUDA buffer with a base-255 header."""
    if not (dist.is_available() and dist.is_initialized()):
        return [data]
    world = get_world_size()
    enc = pickle.dumps(data)
    if len(enc) + 4 > max_size:
        raise ValueError("encoded data exceeds max_size: {}".format(len(enc) + 4))
    dev = "cuda" if dist.get_backend(group) == "nccl" else "cpu"
    slot = torch.zeros(max_size, dtype=torch.uint8)
    slot[:4] = torch.tensor(list(len(enc).to_bytes(4, "little")), dtype=torch.uint8)
    slot[4:4 + len(enc)] = torch.frombuffer(bytearray(enc), dtype=torch.uint8)
    slot = slot.to(dev)
    out = [torch.empty_like(slot) for _ in range(world)]
    dist.all_gather(out, slot, group=group)
    result = []
    try:
        for t in out:
            b = t.cpu().numpy().tobytes()
            n = int.from_bytes(b[:4], "little")
            if n > 0:
                result.append(pickle.loads(b[4:4 + n]))
    except pickle.UnpicklingError:
        raise Exception("Unable to unpickle data from other workers: all ranks must enter all_gather_list together; "
                        "a rank that ran out of memory or finished its epoch early has fallen out of sync.")
    return result


def find_free_port() -> int:
    with socket.socket(socket.AF_INET, socket.SOCK_STREAM) as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]
