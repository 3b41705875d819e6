"""Flat parameter / gradient arena.

Every trainable tensor of the model becomes a view into ONE contiguous fp32 master buffer, its gradient a
view into ONE contiguous fp32 gradient buffer (and, in bf16 mode, the tensor the model computes with is a
view into ONE bf16 mirror).  Consequences:

* gradient scaling, the global-norm reduction, the optimizer update and ``zero_grad`` are each a single
  kernel over the arena (the reference does one launch per tensor per pointwise op: hetseq/optim.py:59-80,
  162-231);
* buckets for the data-parallel reduction are plain slices of the gradient buffer — gradients are written
  into them directly by autograd / by the native backward kernels, there is no copy-in/copy-out as in
  DDP's Reducer (hetseq/controller.py:79-87);
* tensors that a kernel wants fused (query/key/value weights) can be declared *contiguous groups* and then
  be addressed as one packed matrix without changing their state-dict names.

Offsets are aligned to 64 elements (256 bytes) so that every view is a legal TMA base address.
"""
from __future__ import annotations

from typing import Dict, Iterable, List, Optional, Sequence

import torch

ALIGN = 64  # elements


def _round_up(n: int, a: int = ALIGN) -> int:
    return (n + a - 1) // a * a


class ParamArena:
    def __init__(self, params: Sequence[torch.nn.Parameter], groups: Optional[Iterable[Sequence[torch.nn.Parameter]]] = None,
                 device=None, compute_dtype: torch.dtype = torch.float32, bucket_bytes: int = 25 * 1024 * 1024,
                 grad_alloc=None, alloc=None):
        params = [p for p in params if p.requires_grad]
        seen = set()
        uniq = []
        for p in params:
            if id(p) not in seen:
                seen.add(id(p))
                uniq.append(p)
        self.device = torch.device(device) if device is not None else uniq[0].device
        self.compute_dtype = compute_dtype

        # layout order: declared groups first (members adjacent, no padding inside when sizes allow), then the rest
        order: List[torch.nn.Parameter] = []
        group_of: Dict[int, int] = {}
        glist = [list(g) for g in (groups or [])]
        for gi, g in enumerate(glist):
            for p in g:
                if id(p) in seen and id(p) not in group_of:
                    group_of[id(p)] = gi
        emitted = set()
        for p in uniq:
            if id(p) in emitted:
                continue
            if id(p) in group_of:
                for q in glist[group_of[id(p)]]:
                    if id(q) not in emitted and id(q) in seen:
                        order.append(q)
                        emitted.add(id(q))
            else:
                order.append(p)
                emitted.add(id(p))
        self.params = order

        self.offsets: Dict[int, int] = {}
        off = 0
        for p in order:
            self.offsets[id(p)] = off
            off += _round_up(p.numel())
        self.numel = off
        # ``alloc(kind, numel, dtype)`` lets the reducer place an arena in memory it can share with the peers (symmetric
        # buffers of the communication runtime): kind is 'grad', 'master' or 'work'; None -> ordinary device memory
        def _place(kind, numel, dtype):
            t = alloc(kind, numel, dtype) if alloc is not None else None
            if t is None:
                return torch.zeros(numel, dtype=dtype, device=self.device)
            t.zero_()
            return t
        self.master = _place("master", off, torch.float32)
        # 64 spare floats behind the gradients: the per-step statistics vector lives there so that it can ride the
        # same peer-memory mapping / reducer as the gradient buckets
        # (``grad_alloc(numel)`` lets the reducer supply memory it can share with peers, e.g. a symmetric allocation)
        self._grad_full = None
        if alloc is not None:
            self._grad_full = alloc("grad", off + ALIGN, torch.float32)
        elif grad_alloc is not None:
            self._grad_full = grad_alloc(off + ALIGN)
        if self._grad_full is None:
            self._grad_full = torch.zeros(off + ALIGN, dtype=torch.float32, device=self.device)
        else:
            self._grad_full.zero_()
        self.grad = self._grad_full[:off]
        self.tail = self._grad_full[off:]
        self.work = _place("work", off, torch.bfloat16) if compute_dtype == torch.bfloat16 else None

        with torch.no_grad():
            for p in order:
                o, n = self.offsets[id(p)], p.numel()
                self.master[o:o + n].copy_(p.detach().reshape(-1).to(device=self.device, dtype=torch.float32))
                mview = self.master[o:o + n].view(p.shape)
                gview = self.grad[o:o + n].view(p.shape)
                if self.work is not None:
                    self.work[o:o + n].copy_(self.master[o:o + n])
                    p.data = self.work[o:o + n].view(p.shape)
                    p.grad = None
                else:
                    p.data = mview
                    p.grad = gview
                p.master = mview      # fp32 master weights (same tensor as p.data in fp32 mode)
                p.main_grad = gview   # fp32 gradient accumulator every backward path adds into
        self._make_buckets(bucket_bytes)

    # ------------------------------------------------------------------------------------------------
    def _make_buckets(self, bucket_bytes: int):
        """Buckets = contiguous element ranges of the arena, in *reverse* layout order (gradients of the last
        layers are produced first), each about ``bucket_bytes`` of fp32."""
        cap = max(ALIGN, bucket_bytes // 4)
        self.buckets: List[tuple] = []
        self.bucket_of: Dict[int, int] = {}
        end = self.numel
        cur_start = end
        members: List[torch.nn.Parameter] = []
        for p in reversed(self.params):
            o = self.offsets[id(p)]
            if end - o > cap and members:
                self.buckets.append((cur_start, end, members))
                end = cur_start
                members = []
            members.append(p)
            cur_start = o
        if members:
            self.buckets.append((cur_start, end, members))
        for bi, (_, _, ms) in enumerate(self.buckets):
            for p in ms:
                self.bucket_of[id(p)] = bi

    # ------------------------------------------------------------------------------------------------
    def packed(self, plist: Sequence[torch.nn.Parameter], kind: str = "data") -> torch.Tensor:
        """A single 2-D/1-D view covering adjacent parameters (e.g. query/key/value weights -> [3H, H]).
        kind: 'data' (compute tensor), 'master' or 'grad'."""
        first = plist[0]
        o0 = self.offsets[id(first)]
        expect = o0
        for p in plist:
            if self.offsets[id(p)] != expect:
                raise ValueError("parameters are not adjacent in the arena (declare them as a group, sizes must be multiples of %d)" % ALIGN)
            expect += p.numel()
            if p.numel() % ALIGN and p is not plist[-1]:
                raise ValueError("packed parameter sizes must be multiples of %d elements" % ALIGN)
        total = expect - o0
        buf = {"data": self.work if self.work is not None else self.master, "master": self.master, "grad": self.grad}[kind]
        flat = buf[o0:o0 + total]
        if first.dim() == 2:
            return flat.view(-1, first.shape[1])
        return flat

    def zero_grad(self):
        self.grad.zero_()

    def sync_work_from_master(self):
        if self.work is not None:
            from ..ops import kernels
            kernels.cast_to_bf16(self.master, self.work)

    def rebind(self):
        """Re-point ``p.data`` / ``p.grad`` at the arena (after something replaced them, e.g. load_state_dict
        keeps storage, but ``p.grad = None`` by user code does not)."""
        for p in self.params:
            o, n = self.offsets[id(p)], p.numel()
            if self.work is None and p.grad is None:
                p.grad = self.grad[o:o + n].view(p.shape)

    def copy_params_to_master(self):
        """After an external in-place change of ``p.data`` in bf16 mode (state-dict load), refresh the master."""
        with torch.no_grad():
            for p in self.params:
                if self.work is not None:
                    p.master.copy_(p.data.float())

    def state_bytes(self) -> int:
        return self.numel * 4
