"""NVLink peer-to-peer gradient reducer (sm_100a kernels in ``csrc/p2p_allreduce.cu``).

Set-up (once): every rank exports CUDA-IPC handles of its gradient arena and of a small flag array, the handles
are exchanged through ``torch.distributed`` (control plane only), and every rank maps all peers' buffers.  After
that a bucket reduction is ONE kernel launch per rank on a side stream — no NCCL, no host synchronisation, no
copy-in/copy-out (the bucket *is* a slice of the arena the backward kernels wrote into).

Replaces the DDP Reducer + ``ncclAllReduce`` of the reference (hetseq/controller.py:75-90).  Single NVSwitch
domain only (<= 8 ranks on one node); anything else falls back to ``TorchReducer`` (see ``build_reducer``).
"""
from __future__ import annotations

import ctypes
from typing import List

import torch
import torch.distributed as dist

from ..ops import native
from .reducer import BaseReducer

MAX_BUCKETS = 512
MAX_RANKS = 8
ONE_SHOT_BYTES = 256 * 1024
NVLS_MIN_BYTES = 1 << 20          # below this the two barriers dominate; peer loads are as fast
SQ_SLOT = 8                       # arena.tail[SQ_SLOT] accumulates this rank's share of sum(g^2)

_symm = {"handle": None, "tensor": None}


def symmetric_grad_alloc(numel: int):
    """Try to place the gradient arena in torch symmetric memory so that an NVLS multicast mapping exists
    (``multimem.ld_reduce`` / ``multimem.st``).  Returns None when unavailable; the reducer then maps peers with
    plain CUDA IPC and uses the peer load/store kernel only."""
    import os
    if os.environ.get("B200_NO_NVLS"):
        return None
    try:
        import torch.distributed._symmetric_memory as symm_mem
        t = symm_mem.empty(numel, dtype=torch.float32, device=torch.device("cuda", torch.cuda.current_device()))
        hdl = symm_mem.rendezvous(t, dist.group.WORLD)
        _symm["handle"], _symm["tensor"] = hdl, t
        return t
    except Exception as e:  # no fabric / multicast support in this environment
        from ..distributed_utils import log_all
        if dist.get_rank() == 0:
            log_all("| symmetric memory unavailable ({}: {}); NVLS path disabled".format(type(e).__name__, str(e)[:120]))
        _symm["handle"] = _symm["tensor"] = None
        return None


def _declare(lib):
    P, LL, I, U = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_uint
    lib.b200_ipc_export.argtypes = [P, ctypes.c_char_p, ctypes.POINTER(LL)]
    lib.b200_ipc_export.restype = I
    lib.b200_ipc_open.argtypes = [ctypes.c_char_p, LL, ctypes.POINTER(P), ctypes.POINTER(P)]
    lib.b200_ipc_open.restype = I
    lib.b200_ipc_close.argtypes = [P]
    lib.b200_ipc_close.restype = I
    lib.b200_p2p_allreduce.argtypes = [ctypes.POINTER(P), ctypes.POINTER(P), P, I, I, LL, LL, I, I, P, I, I, P, P, P]
    lib.b200_p2p_allreduce.restype = I


class P2PReducer(BaseReducer):
    def __init__(self, arena, world_size: int, blocks: int = 32, tail_blocks: int = 128):
        super().__init__(arena, world_size)
        if arena.device.type != "cuda":
            raise RuntimeError("P2PReducer needs CUDA")
        if world_size > MAX_RANKS:
            raise RuntimeError("P2PReducer supports at most %d ranks (one NVSwitch domain)" % MAX_RANKS)
        if len(arena.buckets) + 1 > MAX_BUCKETS:
            raise RuntimeError("too many buckets for the flag array")
        self.lib = native.cuda_lib()
        _declare(self.lib)
        self.rank = dist.get_rank()
        import os
        self.blocks = int(os.environ.get("B200_P2P_BLOCKS", blocks))            # while backward is still running
        self.tail_blocks = int(os.environ.get("B200_P2P_TAIL_BLOCKS", tail_blocks))  # after backward has finished
        # CTA width of the launches that overlap backward (see b200_p2p_allreduce: threads * 1000 + blocks)
        self.threads = int(os.environ.get("B200_P2P_THREADS", 512))
        self.no_overlap = bool(os.environ.get("B200_P2P_NO_OVERLAP"))
        self._dry = bool(os.environ.get("B200_P2P_DRY"))
        # B200_P2P_DEBUG=1: every bucket is also reduced with torch.distributed (NCCL) from a copy taken right before
        # the launch, and compared after the step; the per-bucket epoch counters are checked against the host's count.
        # Not usable under CUDA-graph capture (it synchronises).
        self.debug = bool(os.environ.get("B200_P2P_DEBUG"))
        self._debug_refs = {}
        self.debug_checked = 0
        dev = arena.device
        self.flags = torch.zeros(2 * MAX_BUCKETS * MAX_RANKS, dtype=torch.int32, device=dev)
        self.counter = torch.zeros(MAX_BUCKETS, dtype=torch.int32, device=dev)
        self.epochs = torch.zeros(MAX_BUCKETS, dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)

        def export(t: torch.Tensor):
            buf = ctypes.create_string_buffer(64)
            off = ctypes.c_longlong(0)
            native.check(self.lib.b200_ipc_export(t.data_ptr(), buf, ctypes.byref(off)), "b200_ipc_export")
            return bytes(buf.raw), int(off.value)

        # gradients in symmetric memory are already mapped by torch's rendezvous (VMM memory has no legacy IPC
        # handle); otherwise export the cudaMalloc'ed arena ourselves
        hdl = _symm["handle"]
        use_symm = (hdl is not None and _symm["tensor"] is not None
                    and _symm["tensor"].data_ptr() == arena._grad_full.data_ptr())
        symm_ptrs = None
        if use_symm:
            symm_ptrs = [int(p) for p in hdl.buffer_ptrs]
            own_off = arena._grad_full.data_ptr() - symm_ptrs[self.rank]
            symm_ptrs = [p + own_off for p in symm_ptrs]
        mine = {"rank": self.rank, "grad": None if use_symm else export(arena.grad), "flags": export(self.flags),
                "numel": arena.numel, "host": _hostname()}
        gathered: List[dict] = [None] * world_size
        dist.all_gather_object(gathered, mine)
        if any(g["numel"] != arena.numel for g in gathered):
            raise RuntimeError("ranks disagree on the arena layout")
        if any(g["host"] != mine["host"] for g in gathered):
            raise RuntimeError("ranks span several hosts: peer memory is unavailable")
        self._opened = []
        data_ptrs = (ctypes.c_void_p * MAX_RANKS)()
        flag_ptrs = (ctypes.c_void_p * MAX_RANKS)()
        for g in gathered:
            r = g["rank"]
            if r == self.rank:
                data_ptrs[r] = arena.grad.data_ptr()
                flag_ptrs[r] = self.flags.data_ptr()
                continue
            for key, arr in (("grad", data_ptrs), ("flags", flag_ptrs)):
                if key == "grad" and use_symm:
                    arr[r] = symm_ptrs[r]
                    continue
                handle, off = g[key]
                p, base = ctypes.c_void_p(), ctypes.c_void_p()
                native.check(self.lib.b200_ipc_open(handle, off, ctypes.byref(p), ctypes.byref(base)), "b200_ipc_open")
                arr[r] = p.value
                self._opened.append(base.value)
        self.data_ptrs, self.flag_ptrs = data_ptrs, flag_ptrs
        # NVLS: multicast address of the arena when it lives in symmetric memory
        self.mc_base = 0
        if use_symm:
            try:
                mc = int(hdl.multicast_ptr or 0)
                self.mc_base = mc + own_off if mc else 0
            except Exception:
                self.mc_base = 0
        self.fused_norm = True            # bucket kernels add sum(g^2) of their slice into arena.tail[SQ_SLOT]
        self.sq_ptr = arena.tail[SQ_SLOT:SQ_SLOT + 1].data_ptr()
        if self.rank == 0:
            from ..distributed_utils import log_all
            log_all("| p2p reducer: {} ranks, {} buckets, NVLS multicast {}".format(
                world_size, len(arena.buckets), "on" if self.mc_base else "off"))
        self.epoch = 0
        self._side = torch.cuda.Stream(device=dev)
        dist.barrier()

    def begin(self, sync_now: bool):
        super().begin(sync_now)
        if self._active:
            self.epoch += 1

    def _launch(self, bi: int):
        self._launched.add(bi)
        s, e, _ = self.arena.buckets[bi]
        n = e - s
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self._side.wait_event(ev)
        if self.debug:
            ref = self.arena.grad[s:e].clone()
            dist.all_reduce(ref)
            self._debug_refs[bi] = ref
        if self._dry:                 # experiment: flag protocol and SM occupancy only, no data movement (WRONG results)
            n = 0
        one_shot = 1 if n * 4 <= ONE_SHOT_BYTES else 0
        mc = self.mc_base if (self.mc_base and n * 4 >= NVLS_MIN_BYTES) else 0
        native.check(self.lib.b200_p2p_allreduce(self.data_ptrs, self.flag_ptrs, self.counter.data_ptr(), self.rank,
                                                 self.world_size, s, n, bi, MAX_BUCKETS, self.epochs.data_ptr(), one_shot,
                                                 self.tail_blocks if self._after_backward else self._overlap_grid(),
                                                 self.sq_ptr, mc, self._side.cuda_stream), "b200_p2p_allreduce")

    def _overlap_grid(self) -> int:
        return self.blocks if self.threads >= 512 else self.threads * 1000 + self.blocks

    def ready(self, params) -> None:
        if self.no_overlap:          # experiment switch: reduce everything after backward
            return
        super().ready(params)

    def _wait_all(self):
        torch.cuda.current_stream().wait_stream(self._side)
        if self.debug and self._debug_refs:
            self._debug_verify()

    def _debug_verify(self):
        """Protocol invariants after a step: every launched bucket equals the NCCL sum of the pre-launch copies, all
        replicas hold identical bits, and each bucket's device-side epoch equals the number of steps the host issued."""
        torch.cuda.synchronize(self.arena.device)
        for bi, ref in sorted(self._debug_refs.items()):
            s, e, _ = self.arena.buckets[bi]
            got = self.arena.grad[s:e]
            scale = float(ref.abs().max().clamp_min(1e-30))
            err = float((got - ref).abs().max()) / scale
            if not err < 1e-5:
                raise RuntimeError("p2p reducer: bucket %d differs from the NCCL sum (rel err %.3e)" % (bi, err))
            chk = torch.stack([got.double().sum(), -got.double().sum()])
            dist.all_reduce(chk, op=dist.ReduceOp.MAX)
            if float(chk[0]) != -float(chk[1]):
                raise RuntimeError("p2p reducer: replicas of bucket %d are not bit-identical" % bi)
            ep = int(self.epochs[bi])
            if ep != self.epoch:
                raise RuntimeError("p2p reducer: bucket %d epoch %d, host expected %d" % (bi, ep, self.epoch))
            self.debug_checked += 1
        self._debug_refs = {}

    def reduce_tail(self):
        """One-shot peer reduction of the statistics vector behind the gradients (64 floats), on the training stream."""
        a = self.arena
        native.check(self.lib.b200_p2p_allreduce(self.data_ptrs, self.flag_ptrs, self.counter.data_ptr(), self.rank,
                                                 self.world_size, a.numel, a.tail.numel(), len(a.buckets), MAX_BUCKETS,
                                                 self.epochs.data_ptr(), 1, 1, None, None,
                                                 torch.cuda.current_stream().cuda_stream), "b200_p2p_allreduce(tail)")

    def all_reduce_flat(self):
        """Reduce the whole arena now (used by tests / when nothing was announced)."""
        self.begin(True)
        self.finish()

    def close(self):
        for base in self._opened:
            self.lib.b200_ipc_close(base)
        self._opened = []


def _hostname() -> str:
    import socket
    return socket.gethostname()
