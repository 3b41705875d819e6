"""NVLink peer-to-peer gradient reducer and sharded optimizer step (sm_100a kernels in ``csrc/p2p_allreduce.cu``
on top of the communication runtime in ``csrc/comm_runtime.cu``).

Set-up (once): the runtime (``CommRuntime``) rendezvouses the ranks over Unix sockets and hands out *symmetric
buffers* — one CUDA virtual-memory allocation per rank that every rank maps, plus an NVLS multicast mapping where
the devices offer one.  The gradient arena, the flag words and (sharded mode) the parameter arena live in such
buffers.  After that a bucket reduction is ONE kernel launch per rank on a side stream — no NCCL, no host
synchronisation, no copy-in/copy-out (the bucket *is* a slice of the arena the backward kernels wrote into).

Two modes:

* **sharded** (default with Adam): per bucket a reduce-scatter kernel leaves the sum of this rank's 1/W slice in
  the local arena; after the statistics tail (loss sums, sum g^2) has been all-reduced, one kernel per rank applies
  scale / clip / Adam to the slices it owns and writes the new parameters into every rank (``multimem.st`` or peer
  stores).  Optimizer work and optimizer traffic are 1/W per rank; gradients cross NVLink once instead of twice.
* **all-reduce**: reduce-scatter + all-gather of the gradients in one kernel, optimizer replicated (used for
  Adadelta, and when the runtime is unavailable and the arena is mapped through torch symmetric memory or CUDA IPC).

Replaces the DDP Reducer + ``ncclAllReduce`` + per-tensor optimizer of the reference (hetseq/controller.py:75-90,
338-347; hetseq/optim.py:162-231).  Single NVSwitch domain only (<= 8 ranks on one node); anything else falls back
to ``TorchReducer`` (see ``build_reducer``).
"""
from __future__ import annotations

import ctypes
import os
import time
from typing import Dict, List, Optional

import torch
import torch.distributed as dist

from ..ops import native
from .reducer import BaseReducer

MAX_BUCKETS = 512
MAX_RANKS = 8
ONE_SHOT_BYTES = 256 * 1024       # 32 blocks of 512 threads; b200_p2p_allreduce refuses more than 64 (co-residency)
NVLS_MIN_BYTES = 1 << 20          # below this the two barriers dominate; peer loads are as fast
SQ_SLOT = 8                       # arena.tail[SQ_SLOT] accumulates this rank's share of sum(g^2)

_symm = {"handle": None, "tensor": None}


class CommBuffer(ctypes.Structure):
    _fields_ = [("handle", ctypes.c_void_p), ("bytes", ctypes.c_longlong), ("local", ctypes.c_void_p),
                ("peer", ctypes.c_void_p * 8), ("multicast", ctypes.c_void_p)]


class _DevMem:
    """Adapter that lets torch wrap a raw device address (``__cuda_array_interface__``)."""

    def __init__(self, ptr: int, numel: int, typestr: str, owner):
        self.__cuda_array_interface__ = {"shape": (numel,), "typestr": typestr, "data": (ptr, False), "version": 2}
        self._owner = owner


class SymBuffer:
    """One symmetric buffer: same size on every rank, every rank's copy mapped here."""

    def __init__(self, rt: "CommRuntime", raw: CommBuffer):
        self.rt = rt
        self.handle = raw.handle
        self.bytes = int(raw.bytes)
        self.local = int(raw.local)
        self.peers = [int(raw.peer[r] or 0) for r in range(rt.world)]
        self.mc = int(raw.multicast or 0)

    def tensor(self, numel: int, dtype: torch.dtype, byte_offset: int = 0) -> torch.Tensor:
        code = {torch.float32: "<f4", torch.int32: "<i4", torch.bfloat16: "<i2", torch.int64: "<i8"}[dtype]
        assert byte_offset + numel * torch.empty((), dtype=dtype).element_size() <= self.bytes
        t = torch.as_tensor(_DevMem(self.local + byte_offset, numel, code, self), device=torch.device("cuda", self.rt.device))
        if dtype == torch.bfloat16:
            t = t.view(torch.bfloat16)
        return t


class CommRuntime:
    """ctypes face of ``csrc/comm_runtime.cu`` (one per process)."""

    _instance: Optional["CommRuntime"] = None
    _failed: Optional[str] = None

    def __init__(self, rank: int, world: int, tag: str):
        self.lib = native.cuda_lib()
        _declare(self.lib)
        self.rank, self.world = rank, world
        self.device = torch.cuda.current_device()
        h = ctypes.c_void_p()
        timeout_ms = int(float(os.environ.get("B200_COMM_TIMEOUT_S", "120")) * 1000)
        native.check(self.lib.b200_comm_init(rank, world, tag.encode(), timeout_ms, ctypes.byref(h)), "b200_comm_init")
        self.handle = h
        self.multicast = bool(self.lib.b200_comm_multicast_supported(h))
        self.by_ptr: Dict[int, SymBuffer] = {}

    @classmethod
    def get(cls) -> Optional["CommRuntime"]:
        """The process-wide runtime, created collectively on first use; None (on every rank) when any rank failed."""
        if cls._instance is not None or cls._failed is not None:
            return cls._instance
        rank, world = dist.get_rank(), dist.get_world_size()
        tag = [None]
        if rank == 0:
            tag[0] = "%d.%x" % (os.getpid(), time.time_ns() & 0xFFFFFFFFFF)
        dist.broadcast_object_list(tag, src=0)
        inst, err = None, ""
        hosts = [None] * world
        dist.all_gather_object(hosts, _hostname())
        if os.environ.get("B200_COMM", "own") != "own":
            err = "disabled by B200_COMM"
        elif world > MAX_RANKS or len(set(hosts)) > 1:
            # Unix sockets and peer mappings end at the host: decided from facts every rank sees, so nobody waits for a
            # rendezvous that cannot happen
            err = "ranks span %d hosts / %d ranks: one NVSwitch domain of at most %d ranks is required" % (len(set(hosts)), world, MAX_RANKS)
        else:
            try:
                inst = cls(rank, world, tag[0])
            except Exception as e:  # noqa: BLE001 - any failure means "use the fallback", decided collectively below
                err = "%s: %s" % (type(e).__name__, str(e)[:200])
        if not _all_ranks_ok(inst is not None):
            if inst is not None:
                inst.close()
            cls._failed = err or "a peer rank could not start the runtime"
            from ..distributed_utils import log_all
            if rank == 0:
                log_all("| comm runtime unavailable ({}); falling back".format(cls._failed))
            return None
        cls._instance = inst
        return inst

    def alloc(self, nbytes: int, multicast: bool = True) -> SymBuffer:
        raw = CommBuffer()
        native.check(self.lib.b200_comm_alloc(self.handle, int(nbytes), int(bool(multicast)), ctypes.byref(raw)), "b200_comm_alloc")
        buf = SymBuffer(self, raw)
        self.by_ptr[buf.local] = buf
        return buf

    def barrier(self):
        native.check(self.lib.b200_comm_barrier(self.handle), "b200_comm_barrier")

    def close(self):
        if self.handle:
            self.lib.b200_comm_destroy(self.handle)
            self.handle = None


def _all_ranks_ok(ok: bool) -> bool:
    flag = torch.tensor([1 if ok else 0], dtype=torch.int32, device=torch.device("cuda", torch.cuda.current_device()))
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    return bool(int(flag.item()))


def symmetric_alloc(numel: int, dtype: torch.dtype = torch.float32):
    """Memory every rank can address, for the arenas (collective: all ranks call it in the same order with the same
    size).  Own runtime first; for fp32 gradient arenas torch symmetric memory is the second choice.  Returns None
    when neither is available (the reducer then maps a plain arena with CUDA IPC)."""
    rt = CommRuntime.get()
    if rt is not None:
        buf = rt.alloc(numel * torch.empty((), dtype=dtype).element_size())
        return buf.tensor(numel, dtype)
    return None


def symmetric_grad_alloc(numel: int):
    """Gradient arena placement: own runtime, else torch symmetric memory (NVLS through torch's mapping), else None."""
    t = symmetric_alloc(numel, torch.float32)
    if t is not None:
        return t
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
    if getattr(lib, "_b200_p2p_declared", False):
        return
    P, LL, I, D = ctypes.c_void_p, ctypes.c_longlong, ctypes.c_int, ctypes.c_double
    PP = ctypes.POINTER(P)
    lib.b200_ipc_export.argtypes = [P, ctypes.c_char_p, ctypes.POINTER(LL)]
    lib.b200_ipc_export.restype = I
    lib.b200_ipc_open.argtypes = [ctypes.c_char_p, LL, PP, PP]
    lib.b200_ipc_open.restype = I
    lib.b200_ipc_close.argtypes = [P]
    lib.b200_ipc_close.restype = I
    lib.b200_p2p_allreduce.argtypes = [PP, PP, P, I, I, LL, LL, I, I, P, I, I, P, P, P]
    lib.b200_p2p_allreduce.restype = I
    lib.b200_p2p_reduce_scatter.argtypes = [PP, PP, P, I, I, LL, LL, I, I, P, I, P, P, P]
    lib.b200_p2p_reduce_scatter.restype = I
    lib.b200_p2p_sharded_adam.argtypes = [PP, P, PP, P, P, P, P, P, I, LL, PP, P, I, I, I, I, P, P, I, P, P]
    lib.b200_p2p_sharded_adam.restype = I
    lib.b200_p2p_set_watchdog.argtypes = [D, P]
    lib.b200_p2p_set_watchdog.restype = I
    lib.b200_comm_init.argtypes = [I, I, ctypes.c_char_p, I, PP]
    lib.b200_comm_init.restype = I
    lib.b200_comm_multicast_supported.argtypes = [P]
    lib.b200_comm_multicast_supported.restype = I
    lib.b200_comm_barrier.argtypes = [P]
    lib.b200_comm_barrier.restype = I
    lib.b200_comm_alloc.argtypes = [P, LL, I, ctypes.POINTER(CommBuffer)]
    lib.b200_comm_alloc.restype = I
    lib.b200_comm_free.argtypes = [P, P]
    lib.b200_comm_free.restype = I
    lib.b200_comm_destroy.argtypes = [P]
    lib.b200_comm_destroy.restype = I
    lib._b200_p2p_declared = True


def _ptr_array(vals) -> ctypes.Array:
    arr = (ctypes.c_void_p * MAX_RANKS)()
    for i, v in enumerate(vals):
        arr[i] = v or None
    return arr


class P2PReducer(BaseReducer):
    def __init__(self, arena, world_size: int, blocks: int = 32, tail_blocks: int = 128, sharded: bool = False):
        super().__init__(arena, world_size)
        if arena.device.type != "cuda":
            raise RuntimeError("P2PReducer needs CUDA")
        if world_size > MAX_RANKS:
            raise RuntimeError("P2PReducer supports at most %d ranks (one NVSwitch domain)" % MAX_RANKS)
        if 2 * len(arena.buckets) + 2 > MAX_BUCKETS:
            raise RuntimeError("too many buckets for the flag array")
        self.lib = native.cuda_lib()
        _declare(self.lib)
        self.rank = dist.get_rank()
        # CTAs of a reduction that runs while backward is still running.  32 is the measured optimum for BERT-base (more
        # CTAs take more issue slots from the GEMMs than the shorter reduction gives back).  What a rank has to reduce
        # per step grows with the model and shrinks with the rank count; when it outgrows what 32 CTAs move under
        # backward, the tail of the reductions is exposed instead: BERT-large bf16 on 2 GPUs (670 MB per rank) 18.24 ms
        # with 32, 17.57 ms with 96.  Hence 32 CTAs per 220 MB of per-rank slice, between 32 and 128.
        per_rank_mb = arena.numel * 4 / max(world_size, 1) / 1e6
        auto_blocks = int(min(128, max(blocks, round(blocks * per_rank_mb / 220.0))))
        self.blocks = int(os.environ.get("B200_P2P_BLOCKS", auto_blocks))       # while backward is still running
        self.tail_blocks = int(os.environ.get("B200_P2P_TAIL_BLOCKS", tail_blocks))  # after backward has finished
        # CTA width of the launches that overlap backward (see b200_p2p_allreduce: threads * 1000 + blocks)
        self.threads = int(os.environ.get("B200_P2P_THREADS", 512))
        self.no_overlap = bool(os.environ.get("B200_P2P_NO_OVERLAP"))
        # update kernels that run NEXT TO the forward pass (--pipelined-update): few CTAs — they are NVLink-bound, and
        # every extra resident warp takes issue slots from the persistent GEMM CTA of its SM
        self.pipelined_blocks = int(os.environ.get("B200_PIPELINED_BLOCKS", "148"))
        # Data paths of the sharded mode: NVLS (multimem) or plain peer loads / stores.  In-switch reduction keeps the
        # inbound traffic of a reduce-scatter at S/W and a multicast store keeps the outbound traffic of the broadcast
        # at S/W — but both loop this rank's own share through the switch as well.  Measured (tools/gpu_probe_p2p.py,
        # 150.7 MB arena): 2 GPUs — reduce-scatter 0.25 ms with peer loads vs 0.34 ms through the switch, Adam +
        # broadcast 0.147 vs 0.245 ms; 8 GPUs — all forms within 5 % (0.32 / 0.23 ms).  B200_RS_PEER / B200_BCAST_PEER
        # = 0 / 1 override the choice.
        small = world_size <= 4
        self.rs_peer = _env_flag("B200_RS_PEER", small)
        self.bcast_peer = _env_flag("B200_BCAST_PEER", small)
        self._dry = bool(os.environ.get("B200_P2P_DRY"))
        # B200_P2P_DEBUG=1: every bucket is also reduced with torch.distributed (NCCL) from a copy taken right before
        # the launch, and compared after the step; the per-bucket epoch counters are checked against the host's count.
        # Not usable under CUDA-graph capture (it synchronises).
        self.debug = bool(os.environ.get("B200_P2P_DEBUG"))
        self._debug_refs = {}
        self.debug_checked = 0
        dev = arena.device
        self.counter = torch.zeros(MAX_BUCKETS, dtype=torch.int32, device=dev)
        self.epochs = torch.zeros(MAX_BUCKETS, dtype=torch.int32, device=dev)
        self._opened: List[int] = []
        self.mc_base = 0
        self.param_mc = self.work_mc = 0
        self.param_ptrs = self.work_ptrs = None
        nflags = 2 * MAX_BUCKETS * MAX_RANKS
        rt = CommRuntime._instance
        gbuf = rt.by_ptr.get(arena._grad_full.data_ptr()) if rt is not None else None
        if gbuf is not None:
            # ---- own runtime: arenas are symmetric buffers already, flags get one too -----------------------------
            self.mapping = "vmm"
            fbuf = rt.alloc(nflags * 4, multicast=False)
            self.flags = fbuf.tensor(nflags, torch.int32)
            self.data_ptrs = _ptr_array(gbuf.peers)
            self.flag_ptrs = _ptr_array(fbuf.peers)
            self.mc_base = gbuf.mc
            pbuf = rt.by_ptr.get(arena.master.data_ptr())
            wbuf = rt.by_ptr.get(arena.work.data_ptr()) if arena.work is not None else None
            if sharded and (pbuf is None if arena.work is None else wbuf is None):
                raise RuntimeError("sharded optimizer step needs the parameter arena in symmetric memory")
            if pbuf is not None:
                self.param_ptrs, self.param_mc = _ptr_array(pbuf.peers), pbuf.mc
            if wbuf is not None:
                self.work_ptrs, self.work_mc = _ptr_array(wbuf.peers), wbuf.mc
            rt.barrier()
        else:
            if sharded:
                raise RuntimeError("sharded optimizer step needs the communication runtime")
            self._map_legacy(arena, world_size, nflags)
        self.sharded = bool(sharded)
        self.fused_norm = True            # bucket kernels add sum(g^2) of their slice into arena.tail[SQ_SLOT]
        self.sq_ptr = arena.tail[SQ_SLOT:SQ_SLOT + 1].data_ptr()
        if self.sharded:
            rng = []
            for s, e, _ in arena.buckets:
                rng += [s, e - s]
            self.ranges = torch.tensor(rng, dtype=torch.int64, device=dev)
        self._setup_watchdog()
        if self.rank == 0:
            from ..distributed_utils import log_all
            log_all("| p2p reducer: {} ranks, {} buckets, mapping {}, NVLS multicast {}, optimizer {}".format(
                world_size, len(arena.buckets), self.mapping, "on" if self.mc_base else "off",
                "sharded (reduce-scatter + owned-slice Adam + parameter broadcast)" if self.sharded else "replicated"))
        self.epoch = 0
        self._side = torch.cuda.Stream(device=dev)
        torch.cuda.synchronize(dev)

    # ---- legacy mappings (torch symmetric memory / CUDA IPC): all-reduce mode only ----------------------------------
    def _map_legacy(self, arena, world_size, nflags):
        dev = arena.device
        self.flags = torch.zeros(nflags, dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)

        def export(t: torch.Tensor):
            buf = ctypes.create_string_buffer(64)
            off = ctypes.c_longlong(0)
            native.check(self.lib.b200_ipc_export(t.data_ptr(), buf, ctypes.byref(off)), "b200_ipc_export")
            return bytes(buf.raw), int(off.value)

        # gradients in torch symmetric memory are already mapped by its rendezvous (VMM memory has no legacy IPC
        # handle); otherwise export the cudaMalloc'ed arena ourselves
        hdl = _symm["handle"]
        use_symm = (hdl is not None and _symm["tensor"] is not None
                    and _symm["tensor"].data_ptr() == arena._grad_full.data_ptr())
        self.mapping = "torch-symm" if use_symm else "cuda-ipc"
        symm_ptrs = None
        own_off = 0
        if use_symm:
            symm_ptrs = [int(p) for p in hdl.buffer_ptrs]
            own_off = arena._grad_full.data_ptr() - symm_ptrs[self.rank]
            symm_ptrs = [p + own_off for p in symm_ptrs]
        err = None
        try:
            mine = {"rank": self.rank, "grad": None if use_symm else export(arena.grad), "flags": export(self.flags),
                    "numel": arena.numel, "host": _hostname(), "err": None}
        except Exception as e:  # noqa: BLE001 - reported to every rank through the gather below
            mine = {"rank": self.rank, "grad": None, "flags": None, "numel": arena.numel, "host": _hostname(), "err": str(e)}
        gathered: List[dict] = [None] * world_size
        dist.all_gather_object(gathered, mine)
        if any(g["err"] for g in gathered):
            raise RuntimeError("IPC export failed on a rank: %s" % [g["err"] for g in gathered if g["err"]][0])
        if any(g["numel"] != arena.numel for g in gathered):
            raise RuntimeError("ranks disagree on the arena layout")
        if any(g["host"] != mine["host"] for g in gathered):
            raise RuntimeError("ranks span several hosts: peer memory is unavailable")
        data_ptrs = (ctypes.c_void_p * MAX_RANKS)()
        flag_ptrs = (ctypes.c_void_p * MAX_RANKS)()
        try:
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
        except Exception as e:  # noqa: BLE001
            err = e
        if not _all_ranks_ok(err is None):      # every rank leaves this constructor the same way
            self.close()
            raise RuntimeError("mapping the peers' arenas failed on a rank: %s" % (err or "peer"))
        self.data_ptrs, self.flag_ptrs = data_ptrs, flag_ptrs
        if use_symm:
            try:
                mc = int(hdl.multicast_ptr or 0)
                self.mc_base = mc + own_off if mc else 0
            except Exception:
                self.mc_base = 0

    def _setup_watchdog(self):
        """A rank that never arrives makes the waiting kernels give up after ``B200_P2P_TIMEOUT_S`` seconds (default
        600) instead of hanging: they record the missing rank in a word of mapped host memory and run to completion;
        the controller looks at the word after every step (``check_watchdog``) and stops the run."""
        self.timeout_s = float(os.environ.get("B200_P2P_TIMEOUT_S", "600"))
        self._err_word = torch.zeros(4, dtype=torch.int32).pin_memory()
        dev_ptr = ctypes.c_void_p(self._err_word.data_ptr())   # unified addressing: pinned host memory is device-addressable as is
        native.check(self.lib.b200_p2p_set_watchdog(self.timeout_s, dev_ptr), "b200_p2p_set_watchdog")

    def watchdog_report(self) -> Optional[str]:
        """What the watchdog recorded before it trapped (None if it never fired)."""
        w = int(self._err_word[0]) & 0xFFFFFFFF
        if not (w & 0x80000000):
            return None
        slot, peer = (w >> 8) & 0x7FFFFF, w & 0xFF
        what = ("bucket %d" % slot if slot < len(self.arena.buckets) else
                "the statistics vector" if slot == len(self.arena.buckets) else "the parameter broadcast")
        return "rank %d did not reach the reduction of %s within %.0f s (dead or stalled peer)" % (peer, what, self.timeout_s)

    def check_watchdog(self):
        """Raise if a reduction gave up waiting for a peer (a host read of pinned memory: no device call)."""
        if int(self._err_word[0]) & 0x80000000:
            raise RuntimeError("peer-to-peer watchdog: " + self.watchdog_report())

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
        grid = self.tail_blocks if self._after_backward else self._overlap_grid()
        mc = self.mc_base if (self.mc_base and n * 4 >= NVLS_MIN_BYTES) else 0
        if self.sharded:
            if self.rs_peer:
                mc = 0
            native.check(self.lib.b200_p2p_reduce_scatter(self.data_ptrs, self.flag_ptrs, self.counter.data_ptr(), self.rank,
                                                          self.world_size, s, n, bi, MAX_BUCKETS, self.epochs.data_ptr(), grid,
                                                          self.sq_ptr, mc, self._side.cuda_stream), "b200_p2p_reduce_scatter")
            return
        one_shot = 1 if n * 4 <= ONE_SHOT_BYTES else 0
        native.check(self.lib.b200_p2p_allreduce(self.data_ptrs, self.flag_ptrs, self.counter.data_ptr(), self.rank,
                                                 self.world_size, s, n, bi, MAX_BUCKETS, self.epochs.data_ptr(), one_shot,
                                                 grid, self.sq_ptr, mc, self._side.cuda_stream), "b200_p2p_allreduce")

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
        """Protocol invariants after a step: every launched bucket equals the NCCL sum of the pre-launch copies (in
        sharded mode: on the slice this rank owns), all replicas hold identical bits, and each bucket's device-side
        epoch equals the number of steps the host issued."""
        torch.cuda.synchronize(self.arena.device)
        for bi, ref in sorted(self._debug_refs.items()):
            s, e, _ = self.arena.buckets[bi]
            got = self.arena.grad[s:e]
            if self.sharded:
                lo, hi = self.slice_of(bi, self.rank)
                got, ref = got[lo - s:hi - s], ref[lo - s:hi - s]
            scale = float(ref.abs().max().clamp_min(1e-30)) if ref.numel() else 1.0
            err = float((got - ref).abs().max()) / scale if ref.numel() else 0.0
            if not err < 1e-5:
                raise RuntimeError("p2p reducer: bucket %d differs from the NCCL sum (rel err %.3e)" % (bi, err))
            if not self.sharded:
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

    # ---- sharded optimizer step -----------------------------------------------------------------------------------
    def slice_of(self, bi: int, rank: int):
        """Element range of bucket ``bi`` that ``rank`` owns (same rule as the kernels: float4 granules, ceil split)."""
        s, e, _ = self.arena.buckets[bi]
        n4 = (e - s) // 4
        per = (n4 + self.world_size - 1) // self.world_size
        lo = min(per * rank, n4)
        hi = min(lo + per, n4)
        return s + 4 * lo, s + 4 * hi

    def _sharded_launch(self, adam_args, m, v, first: int, count: int, slot: int, gate, stream, blocks: int = 0):
        a = self.arena
        bf16 = a.work is not None
        native.check(self.lib.b200_p2p_sharded_adam(
            self.param_ptrs if self.param_ptrs is not None else _ptr_array([a.master.data_ptr() if r == self.rank else 0 for r in range(self.world_size)]),
            (self.param_mc or None) if not (bf16 or self.bcast_peer) else None,
            self.work_ptrs if bf16 else None, (self.work_mc or None) if (bf16 and not self.bcast_peer) else None,
            a.grad.data_ptr(), m.data_ptr(), v.data_ptr(), self.ranges.data_ptr() + 16 * first, count, a.numel,
            self.flag_ptrs, self.counter.data_ptr(), self.rank, self.world_size, slot, MAX_BUCKETS,
            self.epochs.data_ptr(), ctypes.byref(adam_args), int(os.environ.get("B200_SHARDED_ADAM_BLOCKS", "0")) or blocks,
            gate.data_ptr() if gate is not None else None, stream), "b200_p2p_sharded_adam")

    def sharded_adam(self, adam_args, m: torch.Tensor, v: torch.Tensor, gate=None):
        """Scale / clip / Adam on the slices this rank owns + parameter broadcast + zero_grad, one kernel on the
        training stream."""
        self._sharded_launch(adam_args, m, v, 0, len(self.arena.buckets), len(self.arena.buckets) + 1, gate,
                             torch.cuda.current_stream().cuda_stream)

    def sharded_adam_pipelined(self, adam_args, m: torch.Tensor, v: torch.Tensor, gate=None):
        """The same update issued bucket by bucket on the side stream, first layers first: the training stream does not
        wait here — the forward pass that follows calls ``need(params)`` per block and waits only for the bucket that
        holds those parameters, so the parameter broadcast of the later layers runs under the first layers' compute."""
        nb = len(self.arena.buckets)
        ev = torch.cuda.Event()
        ev.record(torch.cuda.current_stream())
        self._side.wait_event(ev)
        self._param_events = {}
        for bi in range(nb - 1, -1, -1):             # buckets are numbered from the END of the arena: last = embeddings
            self._sharded_launch(adam_args, m, v, bi, 1, nb + 1 + bi, gate, self._side.cuda_stream, blocks=self.pipelined_blocks)
            e = torch.cuda.Event()
            e.record(self._side)
            self._param_events[bi] = e

    def need(self, params) -> None:
        """Forward is about to read these parameters: wait for the pipelined update of their buckets (once each)."""
        evs = getattr(self, "_param_events", None)
        if not evs:
            return
        cur = torch.cuda.current_stream()
        for p in params:
            bi = self.arena.bucket_of.get(id(p))
            e = evs.pop(bi, None) if bi is not None else None
            if e is not None:
                cur.wait_event(e)

    def need_all(self) -> None:
        evs = getattr(self, "_param_events", None)
        if evs:
            cur = torch.cuda.current_stream()
            for e in evs.values():
                cur.wait_event(e)
            evs.clear()

    def gather_sharded(self, flats) -> None:
        """Make flat fp32 tensors whose slices are only current on their owners (optimizer moments; the fp32 masters in
        bf16 mode) whole on every rank — checkpointing and state export need that.  Control-plane collectives."""
        if not self.sharded:
            return
        torch.cuda.synchronize(self.arena.device)
        for t in flats:
            for bi in range(len(self.arena.buckets)):
                for r in range(self.world_size):
                    lo, hi = self.slice_of(bi, r)
                    if hi > lo:
                        dist.broadcast(t[lo:hi], src=r)
        torch.cuda.synchronize(self.arena.device)

    def all_reduce_flat(self):
        """Reduce the whole arena now (used by tests / when nothing was announced)."""
        self.begin(True)
        self.finish()

    def close(self):
        for base in self._opened:
            self.lib.b200_ipc_close(base)
        self._opened = []


def _env_flag(name: str, default: bool) -> bool:
    v = os.environ.get(name)
    return default if v is None or v == "" else v not in ("0", "false", "no")


def _hostname() -> str:
    import socket
    return socket.gethostname()
