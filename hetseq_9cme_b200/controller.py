"""The trainer ("Controller").

Public surface of hetseq/controller.py:22-440 — ``train_step``, ``save_checkpoint`` / ``load_checkpoint``,
``get_train_iterator``, ``lr_step`` / ``lr_step_update`` / ``get_lr``, ``get_model`` / ``get_meter`` /
``get_num_updates`` / ``set_num_updates`` / ``zero_grad``, lazy ``model`` / ``optimizer`` / ``lr_scheduler`` — plus a
working ``valid_step`` (the reference stubs validation out, hetseq/train.py:99-100).

What is different underneath:

* no ``DistributedDataParallel``: parameters, gradients and optimizer moments live in flat arenas
  (``parallel/arena.py``); a reducer (``parallel/reducer.py``) sums gradient buckets across ranks while backward
  is still running;
* gradient scale ``W / sample_size`` (after DDP's ``1/W`` average — i.e. ``1 / sample_size`` on the summed
  gradient), global-norm clip, the optimizer update and ``zero_grad`` are ONE kernel driven by device scalars;
* the step never waits for the GPU: loss / grad-norm travel to a pinned mailbox asynchronously and feed the
  meters one step later (``flush_stats`` forces them, e.g. before a log line at the end of an epoch);
* inputs go to the device with non-blocking copies from pinned memory.

Reference quirks that are kept on purpose: ``sample_size`` semantics come from the task (sequence
length for BERT, 1 for MNIST, batch size for NER); without ``--fast-stat-sync`` the logged loss and the
``sample_size`` are those of the last local micro-batch; every rank seeds dropout with ``seed + num_updates``; the
first update runs with ``lr = 0`` when warm-up is enabled.
"""
from __future__ import annotations

import contextlib
import math
import os
from collections import OrderedDict
from itertools import chain
from typing import Any, Dict, List, Optional

import torch

from . import checkpoint_utils, distributed_utils, lr_scheduler as lr_sched, optim, utils
from .meters import AverageMeter, StopwatchMeter, TimeMeter
from .parallel.arena import ParamArena
from .parallel.reducer import build_reducer

_LN2 = math.log(2.0)


class _StatsMailbox:
    """Ring of pinned host slots receiving [loss_sum, nll_sum, sample_size, nsentences, ntokens, gnorm, clipped]
    from the device without blocking the host."""

    FIELDS = 8

    def __init__(self, device, depth: int = 4):
        self.device = device
        self.cuda = device.type == "cuda"
        self.depth = depth
        self.slots = [torch.zeros(self.FIELDS, dtype=torch.float64).pin_memory() if self.cuda
                      else torch.zeros(self.FIELDS, dtype=torch.float64) for _ in range(depth)]
        self.events = [torch.cuda.Event() if self.cuda else None for _ in range(depth)]
        self.pending: List[int] = []
        self.head = 0

    def post(self, dev_vec: torch.Tensor) -> None:
        i = self.head % self.depth
        self.head += 1
        self.slots[i].copy_(dev_vec, non_blocking=True)
        if self.cuda:
            self.events[i].record()
        self.pending.append(i)

    def drain(self, keep_latest: int = 0):
        """Yield completed slots, oldest first, leaving at most ``keep_latest`` in flight."""
        while len(self.pending) > keep_latest:
            i = self.pending.pop(0)
            if self.cuda:
                self.events[i].synchronize()
            yield self.slots[i].clone()


class Controller:
    def __init__(self, args, task, model, criterion=None, dummy_batch=None, oom_batch=None):
        self.args = args
        self.task = task
        self.cuda = torch.cuda.is_available() and not getattr(args, "cpu", False)
        self.device = torch.device("cuda", torch.cuda.current_device()) if self.cuda else torch.device("cpu")
        self._model = model.to(self.device)
        self._dummy_batch = dummy_batch
        self._oom_batch = oom_batch or dummy_batch
        self._lr_scheduler = None
        self._optimizer = None
        self._num_updates = 0
        self._optim_history = None
        self._prev_grad_norm = None
        self._arena: Optional[ParamArena] = None
        self._reducer = None
        self.fast_stat_sync = getattr(args, "fast_stat_sync", False)
        self.world_size = getattr(args, "distributed_world_size", 1) or 1
        self.precision = getattr(args, "precision", "tf32")
        if not self.cuda and self.precision == "bf16":
            self.precision = "fp32"
        if getattr(args, "use_bmuf", False):
            print("| WARNING: --use-bmuf is accepted for compatibility but BMUF is not implemented; "
                  "running synchronous data parallel")
        self._mailbox = _StatsMailbox(self.device)
        self._lazy_stats = self.cuda and not getattr(args, "sync_stats", False)
        self._last_logging_output: Optional[Dict[str, Any]] = None
        # CUDA-graph replay of the fixed-shape step (--cuda-graph)
        self._graph_enabled = bool(getattr(args, "cuda_graph", False)) and self.cuda
        self._graph = None
        self._g_raw = None
        self._graph_cache: Dict[Any, Dict[str, Any]] = {}     # batch signature -> captured step
        self._graph_seen_by_sig: Dict[Any, int] = {}
        self._graph_steps_total = 0
        self._graph_pool = None
        self._g_hyper = None
        self._defer_update = False
        self._pending_hyper = None
        self.init_meters(args)

    # ------------------------------------------------------------------------------------------------
    def init_meters(self, args):
        m = OrderedDict()
        for name in ("train_loss", "train_nll_loss", "valid_loss", "valid_nll_loss"):
            m[name] = AverageMeter()
        m["wps"] = TimeMeter()        # words per second
        m["ups"] = TimeMeter()        # updates per second
        m["wpb"] = AverageMeter()     # words per batch
        m["bsz"] = AverageMeter()     # sentences per batch
        m["gnorm"] = AverageMeter()   # gradient norm
        m["clip"] = AverageMeter()    # fraction of updates clipped
        m["wall"] = TimeMeter()       # wall time in seconds
        m["train_wall"] = StopwatchMeter()
        self.meters = m

    # ---- lazily built engine pieces -----------------------------------------------------------------
    def _build_engine(self):
        model = self._model
        compute_dtype = torch.bfloat16 if self.precision == "bf16" else torch.float32
        params = [p for p in model.parameters() if p.requires_grad]
        groups = model.arena_groups() if hasattr(model, "arena_groups") else None
        bucket_mb = int(getattr(self.args, "bucket_cap_mb", 25))
        alloc = None
        want_sharded = False
        if self.cuda and self.world_size > 1 and getattr(self.args, "comm", "auto") in ("auto", "p2p"):
            from .parallel import p2p
            # sharded optimizer step (reduce-scatter -> Adam on the owned slice -> parameter broadcast) needs the
            # parameters, not only the gradients, in memory that every rank has mapped
            want_sharded = (getattr(self.args, "optimizer", "adam") == "adam"
                            and not getattr(self.args, "no_sharded_optimizer", False)
                            and not os.environ.get("B200_NO_SHARDED")
                            and p2p.CommRuntime.get() is not None)
            bf16 = compute_dtype == torch.bfloat16

            def alloc(kind, numel, dtype):
                if kind == "grad":
                    return p2p.symmetric_grad_alloc(numel)     # NVLS-capable memory when the platform offers it
                if want_sharded and kind == ("work" if bf16 else "master"):
                    return p2p.symmetric_alloc(numel, dtype)
                return None
            # every bucket of the peer-to-peer reducer costs two flag rounds across all ranks plus a launch; measured on
            # 2 B200 (ms/step): 25 MB 8.52, 50 MB 8.43, 100 MB 8.35, 250 MB 8.47, 500 MB 8.73 (profiles/README.md §1)
            bucket_mb = int(getattr(self.args, "p2p_bucket_mb", 100) or bucket_mb)
        bucket_bytes = int(os.environ.get("B200_BUCKET_BYTES") or int(os.environ.get("B200_BUCKET_MB") or bucket_mb) * 1024 * 1024)
        self._arena = ParamArena(params, groups=groups, device=self.device, compute_dtype=compute_dtype,
                                 bucket_bytes=bucket_bytes, alloc=alloc)
        self._reducer = build_reducer(self.args, self._arena, self.world_size, sharded=want_sharded)
        # --precision fp32 on a GPU keeps the native kernels and switches the GEMMs to the 3xTF32 split (ops/gemm.py);
        # --no-native (or --cpu) is what selects plain PyTorch
        use_native = self.cuda and not getattr(self.args, "no_native", False)
        if use_native:
            from .ops import gemm as _gemm
            _gemm.set_strict_fp32(self.precision == "fp32")
        if hasattr(model, "bind_arena"):
            model.bind_arena(self._arena if use_native else None, self._reducer)
        if not use_native and hasattr(model, "force_torch_path"):
            model.force_torch_path = True
        if getattr(self.args, "checkpoint_activations", False):
            model.checkpoint_activations = True          # native blocks recompute; torch path uses torch.utils.checkpoint
        self._nvtx = bool(getattr(self.args, "profile", False)) and self.cuda
        self._hook_handles = []
        if self.world_size > 1:
            # parameters that plain autograd (not a native block) produces gradients for announce themselves
            for p in self._arena.params:
                self._hook_handles.append(p.register_post_accumulate_grad_hook(lambda q: self._reducer.ready([q])))
            self._reducer.broadcast_parameters()
        if self.cuda:
            torch.backends.cuda.matmul.allow_tf32 = self.precision in ("tf32", "bf16")
            torch.backends.cudnn.allow_tf32 = self.precision in ("tf32", "bf16")

    @property
    def arena(self) -> ParamArena:
        if self._arena is None:
            self._build_engine()
        return self._arena

    @property
    def reducer(self):
        if self._arena is None:
            self._build_engine()
        return self._reducer

    @property
    def model(self):
        if self._arena is None:
            self._build_engine()
        return self._model

    @property
    def optimizer(self):
        if self._optimizer is None:
            self._build_optimizer()
        return self._optimizer

    @property
    def lr_scheduler(self):
        if self._lr_scheduler is None:
            self._build_optimizer()
        return self._lr_scheduler

    def _build_optimizer(self):
        arena = self.arena
        self._optimizer = optim.build_optimizer(self.args, arena.params)
        self._optimizer.attach_arena(arena)
        self._lr_scheduler = lr_sched.build_lr_scheduler(self.args, self._optimizer)
        self._lr_scheduler.step_update(0)
        self._init_deferral()

    # ---- checkpoints ---------------------------------------------------------------------------------
    def _meters_state(self):
        return {k: {"type": type(v).__name__, "state": v.state_dict()} for k, v in self.meters.items()}

    def prepare_checkpoint(self):
        """Collective (every rank): with the sharded optimizer step each rank keeps only its slices of the Adam
        moments (and, in bf16 mode, of the fp32 masters) current — gather them so that the master rank can write the
        reference's checkpoint schema unchanged."""
        self.flush_pending_update()
        self.flush_stats()
        red = self._reducer
        if red is not None and getattr(red, "sharded", False) and self._optimizer is not None:
            flats = list(self._optimizer.optimizer.flat_state_tensors())
            if self._arena.work is not None:
                flats.append(self._arena.master)
            red.gather_sharded(flats)

    def save_checkpoint(self, filename, extra_state):
        """Write all training state (master rank only; ``prepare_checkpoint`` must have run on every rank)."""
        self.flush_stats()
        if distributed_utils.is_master(self.args):
            extra_state = dict(extra_state)
            extra_state["train_meters"] = self._meters_state()
            checkpoint_utils.save_state(filename, self.args, self._export_model_state(), None, self.optimizer,
                                        self.lr_scheduler, self.get_num_updates(), self._optim_history, extra_state)

    def _export_model_state(self):
        """fp32 state dict from the master weights (bf16 mode keeps a bf16 working copy in ``p.data``)."""
        sd = self.get_model().state_dict()
        if self._arena is not None and self._arena.work is not None:
            named = dict(self.get_model().named_parameters())
            for k in list(sd.keys()):
                p = named.get(k)
                if p is not None and hasattr(p, "master"):
                    sd[k] = p.master
        return sd

    def load_checkpoint(self, filename, reset_optimizer=False, reset_lr_scheduler=False, optimizer_overrides=None,
                        reset_meters=False):
        """Restore model / optimizer / scheduler / meters; returns the checkpoint's ``extra_state`` (or None)."""
        extra_state, self._optim_history, last_optim_state = None, [], None
        if os.path.exists(filename):
            state = checkpoint_utils.load_checkpoint_to_cpu(filename)
            _ = self.arena   # parameters must already be arena views so that loading writes into the arena
            try:
                self.get_model().load_state_dict(state["model"], strict=True)
            except Exception:
                raise Exception("Cannot load model parameters from checkpoint {}; please ensure that the "
                                "architectures match.".format(filename))
            if self._arena.work is not None:
                named = dict(self.get_model().named_parameters())
                with torch.no_grad():
                    for k, v in state["model"].items():
                        if k in named and hasattr(named[k], "master"):
                            named[k].master.copy_(v.to(self.device, torch.float32))
                self._arena.sync_work_from_master()
            extra_state = state.get("extra_state") or {}
            self._optim_history = state["optimizer_history"]
            last_optim_state = state.get("last_optimizer_state", None)

        if last_optim_state is not None and not reset_optimizer:
            self._build_optimizer()
            last_optim = self._optim_history[-1]
            assert last_optim["optimizer_name"] == self.optimizer.__class__.__name__, \
                "Optimizer does not match; please reset the optimizer (--reset-optimizer)."
            if not reset_lr_scheduler:
                self.lr_scheduler.load_state_dict(last_optim["lr_scheduler_state"])
            self.optimizer.load_state_dict(last_optim_state, optimizer_overrides)
            self.set_num_updates(last_optim["num_updates"])

        if extra_state is not None:
            epoch = extra_state.get("train_iterator", {}).get("epoch", 0)
            print("| loaded checkpoint {} (epoch {} @ {} updates)".format(filename, epoch, self.get_num_updates()))
            self.lr_step(epoch)
            tm = extra_state.pop("train_meters", None)
            if tm is not None and not reset_meters:
                for k, rec in tm.items():
                    if k in self.meters and isinstance(rec, dict) and "state" in rec:
                        self.meters[k].load_state_dict(rec["state"])
                for meter in self.meters.values():
                    if isinstance(meter, TimeMeter):
                        meter.reset()
        else:
            print("| no existing checkpoint found {}".format(filename))
        return extra_state

    # ---- data -----------------------------------------------------------------------------------------
    def get_train_iterator(self, epoch, combine=True, load_dataset=True):
        if load_dataset:
            print("| loading train data for epoch {}".format(epoch))
            self.task.load_dataset(self.args.train_subset)
        return self.task.get_batch_iterator(
            dataset=self.task.dataset(self.args.train_subset), max_tokens=self.args.max_tokens,
            max_sentences=self.args.max_sentences, max_positions=None, ignore_invalid_inputs=True,
            required_batch_size_multiple=self.args.required_batch_size_multiple, seed=self.args.seed,
            num_shards=self.world_size, shard_id=getattr(self.args, "distributed_rank", 0) or 0,
            num_workers=self.args.num_workers, epoch=epoch)

    def _range(self, name):
        """NVTX range for ``--profile`` (data / forward_backward / reduce_wait / update); no-op otherwise."""
        if getattr(self, "_nvtx", False):
            return torch.cuda.nvtx.range(name)
        return contextlib.nullcontext()

    def _prepare_sample(self, sample):
        if utils.sample_is_empty(sample):
            return None
        from .data.packed import PackedBertBatch
        if isinstance(sample, PackedBertBatch):          # --device-batch-assembly: one H2D copy + one kernel
            return self._batch_assembler().assemble(sample)
        if self.cuda:
            sample = utils.move_to_cuda(sample, self.device, non_blocking=True)
        return sample

    def _settle_gc(self, again: bool = False):
        """Move everything built so far (model, arena views, datasets: ~170k objects for BERT-base) into the garbage
        collector's permanent generation.  A full collection over them takes 20-80 ms of host time, and with the
        GPU at most one step behind the host that is a bubble on the device: measured as one 27 ms step in every
        ~30 on B200 (profiles/README.md §8).  Frozen objects are skipped by later collections, which then only scan
        what the steps themselves allocate."""
        if getattr(self, "_gc_settled", False) and not again:
            return
        import gc
        gc.collect()
        gc.freeze()
        self._gc_settled = True

    def _batch_assembler(self):
        if getattr(self, "_assembler", None) is None:
            from .data.packed import DeviceBatchAssembler
            self._assembler = DeviceBatchAssembler(self.device if self.cuda else None)
        return self._assembler

    def _set_seed(self):
        seed = self.args.seed + self.get_num_updates()
        torch.manual_seed(seed)
        if self.cuda:
            torch.cuda.manual_seed(seed)
            from .ops import kernels
            kernels.Workspace.get(self.device).set_seed(seed)

    # ---- the hot loop -----------------------------------------------------------------------------------
    def train_step(self, samples, dummy_batch=False, raise_oom=False):
        """Forward, backward, cross-rank reduction and parameter update for one group of micro-batches.

        Eager launches and CUDA-graph replays issue the SAME sequence of cross-rank operations (one reduction per
        bucket, one reduction of the statistics tail, the parameter broadcast of the sharded update), so ranks may
        take different paths in the same step — a rank with the end-of-epoch filler batch or a new batch shape runs
        eagerly (or captures) while its peers replay."""
        if self._dummy_batch is None:
            self._dummy_batch = samples[0]
        model, optimizer, reducer = self.model, self.optimizer, self.reducer
        self._settle_gc()
        if self._graph_enabled and not dummy_batch and self._graphable(samples):
            return self._graph_train_step(samples)
        self._set_seed()
        model.train()
        if not dummy_batch:
            self.meters["train_wall"].start()
        if self.cuda:
            return self._eager_step_cuda(samples, dummy_batch)
        return self._eager_step_cpu(samples, dummy_batch)

    def _run_micro_batches(self, samples, on_loss):
        """Forward + backward of every micro-batch; reductions start under the last one.  Returns host-side sums
        [sample_size, nsentences, ntokens] and the last real micro-batch's (logging_output, sample_size)."""
        reducer = self._reducer
        host_stats = [0.0, 0.0, 0.0]
        last_logging, last_sample_size = None, 0
        for i, sample in enumerate(samples):
            sample = self._prepare_sample(sample)
            ignore_grad = sample is None
            if ignore_grad:
                # short rank at the end of an epoch: stay in lock-step with a zero-weight batch
                sample = self._prepare_sample(self._dummy_batch)
            reducer.begin(sync_now=(i == len(samples) - 1))
            try:
                with self._range("forward_backward"):
                    loss, sample_size, logging_output = self.task.train_step(sample, self._model, self._optimizer, ignore_grad)
            except RuntimeError as e:
                if "out of memory" in str(e):
                    raise RuntimeError("ran out of memory with exception: {}".format(e))
                raise
            if not ignore_grad:
                host_stats[0] += sample_size
                host_stats[1] += logging_output.get("nsentences", 0)
                host_stats[2] += logging_output.get("ntokens", 0)
                on_loss(logging_output)
                self._note_mlm_overflow()
                last_logging, last_sample_size = logging_output, sample_size
        return host_stats, last_logging, last_sample_size

    def _note_mlm_overflow(self):
        """The labelled-positions MLM head has a fixed capacity; positions beyond it are dropped.  Their count rides
        the statistics mailbox (field 7, no sync) and turns into a warning when the host reads it."""
        cnt = getattr(self._model, "last_mlm_count", None)
        if cnt is not None and self.cuda:
            if getattr(self, "_mlm_over", None) is None:
                self._mlm_over = torch.zeros(1, dtype=torch.float64, device=self.device)
            self._mlm_over += (cnt[1] - cnt[0]).clamp_min(0).double()

    def _device_update(self, tail, sample_size_dev, mult, hyper_dev=None, advance=True):
        """After backward: wait for the bucket reductions, reduce the statistics tail, then scale / clip / update in
        one kernel (on the slices this rank owns when the reducer is sharded).  Runs eagerly or under graph capture."""
        from .ops import kernels
        arena, reducer, opt = self._arena, self._reducer, self._optimizer.optimizer
        ws = kernels.Workspace.get(self.device)
        with self._range("reduce_wait"):
            reducer.finish()
            if self.world_size > 1:
                reducer.reduce_tail()
        if getattr(reducer, "fused_norm", False):
            from .parallel.p2p import SQ_SLOT
            norm_sq = tail[SQ_SLOT:SQ_SLOT + 1]       # summed over ranks by reduce_tail: no separate norm pass
        else:
            kernels.sumsq(arena.grad, ws.norm_sq, ws=ws)
            norm_sq = ws.norm_sq
        max_norm = float(getattr(self.args, "clip_norm", 0.0))
        if self._defer_update:
            # The sharded update ends with the parameter broadcast — NVLink-bound and, placed here, fully exposed.
            # Park what it needs instead ({sum g^2, sample size, valid}, this step's schedule values) and let the NEXT
            # step issue it bucket by bucket next to its forward pass (_apply_pending_update).  gnorm / clip for the
            # log are derived here from the same scalars.
            pend = self._pend
            pend[0:1].copy_(norm_sq)
            if sample_size_dev is not None:
                pend[1:2].copy_(sample_size_dev.float() / mult if mult != 1.0 else sample_size_dev)
            else:
                pend[1:2].fill_(1.0 / mult)
            pend[2:3].fill_(1.0)                  # host side of the parking: _park_host(), by the caller (not under capture)
            scale = torch.where(pend[1:2] > 0, 1.0 / pend[1:2].clamp_min(1e-30), torch.zeros_like(pend[1:2]))
            gnorm = pend[0:1].clamp_min(0).sqrt() * scale
            ws.step_stats[0:1].copy_(gnorm)
            ws.step_stats[2:3].copy_(((gnorm > max_norm) & (max_norm > 0)).float())
            return ws.step_stats
        kw = dict(grad_scale_mult=mult, max_norm=max_norm, norm_sq=norm_sq,
                  sample_size=sample_size_dev, stats_out=ws.step_stats, hyper_dev=hyper_dev, advance=advance)
        with self._range("update"):
            if getattr(reducer, "sharded", False):
                opt.sharded_step(reducer, **kw)
            else:
                opt.fused_step(zero_grad=True, **kw)
        return ws.step_stats

    # ---- deferred (pipelined) sharded update ---------------------------------------------------------------------
    def _init_deferral(self):
        red = self._reducer
        want = bool(getattr(self.args, "pipelined_update", False) or os.environ.get("B200_PIPELINED_UPDATE")) \
            and not os.environ.get("B200_NO_DEFER")
        self._defer_update = bool(self.cuda and getattr(red, "sharded", False) and want)
        self._pending_hyper = None
        if self._defer_update:
            self._pend = torch.zeros(4, dtype=torch.float32, device=self.device)          # sum g^2, sample size, valid
            self._pend_hyper_dev = torch.zeros(4, dtype=torch.float32, device=self.device)
            self._pend_hyper_host = [torch.zeros(4, dtype=torch.float32).pin_memory() for _ in range(4)]
            self._pend_hyper_i = 0

    def _stage_pending_hyper(self):
        """Schedule values of the parked update -> device (async copy from a small pinned ring; before the launch /
        replay that applies it)."""
        if not self._defer_update:
            return
        h = self._pend_hyper_host[self._pend_hyper_i % 4]
        self._pend_hyper_i += 1
        vals = self._pending_hyper or [0.0, 0.0, 1.0, 0.0]
        for i, v in enumerate(vals):
            h[i] = v
        self._pend_hyper_dev.copy_(h, non_blocking=True)

    def _apply_pending_update(self, pipelined=True):
        """Issue the parked update (every step, also when nothing is parked: the kernels carry a flag round that all
        ranks must run; the device-side gate makes them skip the data).  Pipelined: on the reducer's side stream, the
        forward pass waits per bucket.  Otherwise on the training stream (flush)."""
        if not self._defer_update:
            return
        opt, red = self._optimizer.optimizer, self._reducer
        with self._range("update"):
            opt.sharded_step(red, grad_scale_mult=1.0, max_norm=float(getattr(self.args, "clip_norm", 0.0)),
                             norm_sq=self._pend[0:1], sample_size=self._pend[1:2], stats_out=None,
                             hyper_dev=self._pend_hyper_dev, advance=False, gate=self._pend[2:3], pipelined=pipelined)
        if not pipelined:
            self._pend[2:3].zero_()         # same stream as the launch: ordered after it
        elif not (hasattr(self._model, "bind_arena") and not getattr(self._model, "force_torch_path", False)):
            red.need_all()                  # models on plain PyTorch ops do not announce what they read

    def _park_host(self):
        """Host half of parking an update: remember the schedule values (lr, bias-corrected step size, decay) that are
        current NOW — the kernels that apply it run after the scheduler has moved on."""
        if self._defer_update:
            opt = self._optimizer.optimizer
            self._pending_hyper = list(opt.hyper_values(opt._global_step + 1))

    def _note_update_applied(self):
        """Host bookkeeping after the launch / replay that applied a parked update."""
        if self._defer_update and self._pending_hyper is not None:
            self._optimizer.optimizer._global_step += 1
            self._pending_hyper = None

    def flush_pending_update(self):
        """Apply a parked update now (collective: every rank calls it at the same point) — before anything reads the
        parameters outside a training step: validation, checkpoints, the end of training."""
        if not getattr(self, "_defer_update", False) or self._pending_hyper is None:
            return
        self._stage_pending_hyper()
        self._apply_pending_update(pipelined=False)
        self._note_update_applied()

    def _eager_step_cuda(self, samples, dummy_batch):
        dev = self.device
        tail = self._arena.tail
        self._stage_pending_hyper()
        self._apply_pending_update()
        self._note_update_applied()
        tail.zero_()
        stat_sum = torch.zeros(self._mailbox.FIELDS, dtype=torch.float64, device=dev)

        def on_loss(lo):
            stat_sum[0] += _as_dev_scalar(lo.get("loss", 0.0), dev)
            stat_sum[1] += _as_dev_scalar(lo.get("nll_loss", 0.0), dev)
        host_stats, last_logging, last_sample_size = self._run_micro_batches(samples, on_loss)
        if dummy_batch:
            with self._range("reduce_wait"):
                self._reducer.finish()
            if self._defer_update:
                self._pend[2:3].zero_()     # the parked update was applied above and this step parks none
            return None

        # ---- statistics: global sums ride the tail (one peer reduction, no NCCL call, no host sync) --------------
        if self.fast_stat_sync:
            if getattr(self, "_stat_stage", None) is None:
                self._stat_stage = [torch.zeros(3, dtype=torch.float32).pin_memory() for _ in range(4)]
                self._stat_stage_i = 0
            st = self._stat_stage[self._stat_stage_i % 4]
            self._stat_stage_i += 1
            st[0], st[1], st[2] = host_stats[0], host_stats[1], host_stats[2]
            tail[0:2] = stat_sum[0:2].float()
            tail[2:5].copy_(st, non_blocking=True)
            sample_size_dev, mult = tail[2:3], 1.0
        else:
            if self.world_size > 1 and not self._graph_enabled:
                # reference's slow path (hetseq/controller.py:316-329): host statistics + previous grad norm of every
                # rank; with --cuda-graph the replicas' identity is guaranteed by construction and the check is skipped
                gathered = distributed_utils.all_gather_list([host_stats[:3], self._prev_grad_norm_host()])
                norms = [g[1] for g in gathered]
                if not getattr(self.args, "use_bmuf", False):
                    assert all(n == norms[0] for n in norms) or all(n is None or math.isnan(n) or math.isinf(n) for n in norms), \
                        "Fatal error: gradients are inconsistent between workers"
            sample_size_dev = None
            mult = 1.0 / float(last_sample_size) if last_sample_size and last_sample_size > 0 else 1.0
        step_stats = self._device_update(tail, sample_size_dev, mult)
        self._park_host()
        if self.fast_stat_sync:
            stat_sum[0:5] = tail[0:5].double()
        else:
            ll = last_logging or {}
            stat_sum[0] = _as_dev_scalar(ll.get("loss", 0.0), dev)
            stat_sum[1] = _as_dev_scalar(ll.get("nll_loss", 0.0), dev)
            stat_sum[2:5] = torch.tensor([float(last_sample_size), float(ll.get("nsentences", 0)), float(ll.get("ntokens", 0))],
                                         dtype=torch.float64).to(dev, non_blocking=True)
        stat_sum[5:7] = step_stats[0:3:2].double()
        if getattr(self, "_mlm_over", None) is not None:
            stat_sum[7:8] = self._mlm_over
            self._mlm_over.zero_()
        return self._finish_step(stat_sum)

    def _eager_step_cpu(self, samples, dummy_batch):
        dev = self.device
        optimizer, reducer = self._optimizer, self._reducer
        stat_sum = torch.zeros(self._mailbox.FIELDS, dtype=torch.float64, device=dev)

        def on_loss(lo):
            stat_sum[0] += _as_dev_scalar(lo.get("loss", 0.0), dev)
            stat_sum[1] += _as_dev_scalar(lo.get("nll_loss", 0.0), dev)
        host_stats, last_logging, last_sample_size = self._run_micro_batches(samples, on_loss)
        with self._range("reduce_wait"):
            reducer.finish()
        if dummy_batch:
            return None
        if self.fast_stat_sync:
            stat_sum[2:5] += torch.tensor(host_stats[:3], dtype=torch.float64)
            if self.world_size > 1:
                torch.distributed.all_reduce(stat_sum[:5])
            ss = float(stat_sum[2].item())
        else:
            if self.world_size > 1:
                gathered = distributed_utils.all_gather_list([host_stats[:3], self._prev_grad_norm_host()])
                norms = [g[1] for g in gathered]
                if not getattr(self.args, "use_bmuf", False):
                    assert all(n == norms[0] for n in norms) or all(n is None or math.isnan(n) or math.isinf(n) for n in norms), \
                        "Fatal error: gradients are inconsistent between workers"
            ss = float(last_sample_size)
            ll = last_logging or {}
            stat_sum[2:5] = torch.tensor([ss, float(ll.get("nsentences", 0)), float(ll.get("ntokens", 0))], dtype=torch.float64)
            if last_logging is not None:
                stat_sum[0] = _as_dev_scalar(ll.get("loss", 0.0), dev)
                stat_sum[1] = _as_dev_scalar(ll.get("nll_loss", 0.0), dev)
        clip_norm = float(getattr(self.args, "clip_norm", 0.0))
        if ss > 0:
            optimizer.multiply_grads(1.0 / ss)
        gnorm = optimizer.clip_grad_norm(clip_norm)
        optimizer.step()
        optimizer.zero_grad()
        stat_sum[5] = float(gnorm)
        stat_sum[6] = 1.0 if (clip_norm > 0 and float(gnorm) > clip_norm) else 0.0
        return self._finish_step(stat_sum)

    def _finish_step(self, stat_sum):
        if hasattr(self._reducer, "check_watchdog"):
            self._reducer.check_watchdog()          # a peer never arrived: stop here, with its rank in the message
        self.set_num_updates(self.get_num_updates() + 1)
        self.task.update_step(self._num_updates)
        # ---- meters (asynchronously on CUDA) --------------------------------------------------------------
        self._mailbox.post(stat_sum)
        self._consume_stats(keep_latest=1 if self._lazy_stats else 0)
        self.meters["train_wall"].stop()
        return self._last_logging_output if self._last_logging_output is not None else {
            "loss": 0.0, "nll_loss": 0.0, "ntokens": 0, "nsentences": 0, "sample_size": 0}

    # ---- CUDA-graph path ------------------------------------------------------------------------------------
    # The fixed-shape step (forward, backward, bucket reductions, norm, fused update, statistics) is captured once
    # and replayed; everything that changes from step to step lives in device memory that is refreshed before
    # the replay with asynchronous copies from pinned buffers: the batches, the dropout seed, and {lr, bias-corrected
    # step size, decay factor}.  Launch overhead per step drops from ~400 Python->ctypes->CUDA launches per
    # micro-batch to one.  A step of several micro-batches (--update-freq) is ONE graph: gradients accumulate in the
    # arena and the reductions start under the last micro-batch's backward.
    @staticmethod
    def _sig1(sample):
        ts = sample if isinstance(sample, (list, tuple)) else list(sample.values())
        return tuple((tuple(t.shape), str(t.dtype)) for t in ts if torch.is_tensor(t))

    def _sig(self, samples):
        return tuple(self._sig1(s) for s in samples)

    def _graphable(self, samples) -> bool:
        if len(samples) < 1 or len(samples) > self.MAX_GRAPH_MICRO or any(utils.sample_is_empty(x) for x in samples):
            return False
        if not all(isinstance(x, (list, tuple, dict)) for x in samples):
            return False
        if not hasattr(self.task, "_sample_size"):
            return False
        if type(self._reducer).__name__ == "TorchReducer":
            return False       # NCCL / gloo collectives stay outside graphs (decided by flags: the same on every rank)
        sig = self._sig(samples)
        if sig in self._graph_cache:
            return True
        # a few eager steps first (lazy initialisation, warm caches); later shapes are captured on their second visit.
        # Batches padded to the longest sequence (NER) produce one graph per distinct length, up to MAX_GRAPHS.
        seen = self._graph_seen_by_sig.get(sig, 0) + 1
        self._graph_seen_by_sig[sig] = seen
        self._graph_steps_total += 1
        if len(self._graph_cache) >= self.MAX_GRAPHS:
            return False
        return self._graph_steps_total > 3 and (seen > 1 or not self._graph_cache)

    MAX_GRAPHS = 24
    MAX_GRAPH_MICRO = 16

    def _graph_body(self, st):
        """One full step on the static buffers (runs eagerly while capturing)."""
        from .ops import kernels
        arena, reducer = self._arena, self._reducer
        tail = arena.tail
        self._apply_pending_update()          # last step's parked update, next to this step's forward pass
        tail.zero_()
        n = len(st["samples"])
        for i in range(n):
            if st["raw"][i] is not None:      # packed batch: the assembly kernel is the first node of the micro-batch
                kernels.assemble_bert_batch(st["raw"][i], *st["raw_dims"][i], out=st["samples"][i])
            reducer.begin(sync_now=(i == n - 1))
            loss, sample_size, logging_output = self.task.train_step(st["samples"][i], self._model, self.optimizer, False)
            lf = loss.detach().reshape(1).float()
            tail[0:1] += lf
            tail[1:2] += lf
            if i == n - 1:
                st["last_loss"].copy_(lf)
        tail[2:5] += st["const"]
        if self.fast_stat_sync:
            sample_size_dev = tail[2:3]                       # global sum after reduce_tail
        else:
            sample_size_dev = st["const_last"][0:1]           # the reference scales by the last local micro-batch's size
        step_stats = self._device_update(tail, sample_size_dev, 1.0, hyper_dev=self._g_hyper, advance=False)
        if self.fast_stat_sync:
            self._g_out[0:5] = tail[0:5].double()
        else:
            self._g_out[0:2] = st["last_loss"].double().expand(2)
            self._g_out[2:5] = st["const_last"].double()
        self._g_out[5:7] = step_stats[0:3:2].double()

    def _graph_capture(self, samples):
        dev = self.device
        from .data.packed import PackedBertBatch
        from .ops import kernels as _k
        from .ops import native
        sig = self._sig(samples)
        st = {"samples": [], "raw": [], "raw_dims": []}
        ss_sum = ns_sum = ntok_sum = 0.0
        for sample in samples:
            if isinstance(sample, PackedBertBatch):
                raw, dims = sample.raw.to(dev).clone(), sample.dims
                static = _k.assemble_bert_batch(raw, *dims)
            elif isinstance(sample, dict):
                raw, dims = None, None
                static = {k: (v.to(dev).clone() if torch.is_tensor(v) else v) for k, v in sample.items()}
            else:
                raw, dims = None, None
                static = [t.to(dev).clone() for t in sample]
            st["samples"].append(static); st["raw"].append(raw); st["raw_dims"].append(dims)
            ss = float(self.task._sample_size(static))
            nt = float(self.task._num_tokens(static))
            ns = float(self.task._num_sentences(static)) if hasattr(self.task, "_num_sentences") else ss
            ss_sum += ss; ns_sum += ns; ntok_sum += nt
        if self._g_hyper is None:           # shared by every captured shape
            self._g_hyper = torch.zeros(4, dtype=torch.float32, device=dev)
            self._g_hyper_host = torch.zeros(4, dtype=torch.float32).pin_memory()
            self._g_out = torch.zeros(self._mailbox.FIELDS, dtype=torch.float64, device=dev)
            self._graph_pool = torch.cuda.graph_pool_handle()
        st["const"] = torch.tensor([ss_sum, ns_sum, ntok_sum], dtype=torch.float32, device=dev)
        st["const_last"] = torch.tensor([ss, ns, nt], dtype=torch.float32, device=dev)
        st["last_loss"] = torch.zeros(1, dtype=torch.float32, device=dev)
        if not self._defer_update:
            self._stage_hyper()
        torch.cuda.synchronize()
        # capture on a side stream, as torch requires; state changes made while capturing are real work that is
        # NOT executed (capture only records), so the first real execution is the first replay below
        g = torch.cuda.CUDAGraph()
        self._model.train()
        before = native.launch_count()
        # replays are serial, so all shapes share one pool.  thread_local: the loader threads keep pinning host memory
        # while this thread captures — in the default (global) mode such a call from ANY thread invalidates the capture
        with torch.cuda.graph(g, pool=self._graph_pool, capture_error_mode="thread_local"):
            self._graph_body(st)
        st["launches"] = native.launch_count() - before        # kernels of ours inside one replay
        native.add_launch_count(-st["launches"])                # capture recorded them, it did not run them
        st["graph"] = g
        self._graph = g
        self._graph_cache[sig] = st
        self._settle_gc(again=True)
        print("| captured the training step in a CUDA graph (#{}: {} micro-batch(es), sample size {})".format(
            len(self._graph_cache), len(samples), ss_sum))
        return st

    def _stage_hyper(self):
        opt = self.optimizer.optimizer
        vals = opt.hyper_values(opt._global_step + 1)
        for i, v in enumerate(vals):
            self._g_hyper_host[i] = v
        self._g_hyper.copy_(self._g_hyper_host, non_blocking=True)

    def _graph_train_step(self, samples):
        self.meters["train_wall"].start()
        self._set_seed()
        st = self._graph_cache.get(self._sig(samples))
        if st is None:
            st = self._graph_capture(samples)
        self._graph = st["graph"]
        self._g_raw = st["raw"][0]
        for i, sample in enumerate(samples):
            static = st["samples"][i]
            if st["raw"][i] is not None:
                st["raw"][i].copy_(sample[0], non_blocking=True)
            elif isinstance(sample, dict):
                for k, src in sample.items():
                    if torch.is_tensor(src):
                        static[k].copy_(src, non_blocking=True)
            else:
                for dst, src in zip(static, sample):
                    dst.copy_(src, non_blocking=True)
        if self._defer_update:
            self._stage_pending_hyper()
        else:
            self._stage_hyper()
        st["graph"].replay()
        from .ops import native
        native.add_launch_count(st["launches"])
        if self._defer_update:
            self._note_update_applied()        # the replay applied the previous step's update ...
            self._park_host()                  # ... and parked this one's
        else:
            self.optimizer.optimizer._global_step += 1
        return self._finish_step(self._g_out)

    def watchdog_report(self) -> Optional[str]:
        """What the peer-to-peer kernels' watchdog recorded (which rank never arrived), or None.  The word lives in
        mapped host memory: reading it needs no device call."""
        red = self._reducer
        if red is not None and hasattr(red, "watchdog_report"):
            try:
                return red.watchdog_report()
            except Exception:
                return None
        return None

    def _prev_grad_norm_host(self):
        return self._prev_grad_norm

    def _consume_stats(self, keep_latest: int):
        for v in self._mailbox.drain(keep_latest):
            loss, nll, sample_size, nsent, ntok, gnorm, clipped = (float(x) for x in v[:7])
            if float(v[7]) > 0 and not getattr(self, "_warned_mlm_overflow", False):
                self._warned_mlm_overflow = True
                print("| WARNING: {:.0f} labelled MLM positions exceeded the head's capacity (--max-predictions-per-seq "
                      "{}) and were left out of the loss; raise the flag".format(
                          float(v[7]), getattr(self._model, "max_predictions_per_seq", "?")), flush=True)
            if self.fast_stat_sync and sample_size > 0:
                # reference: global sums divided by sample_size and converted to base 2 (controller.py:298-305)
                loss, nll = loss / (sample_size * _LN2), nll / (sample_size * _LN2)
            self._prev_grad_norm = gnorm
            m = self.meters
            m["wps"].update(ntok)
            m["ups"].update(1.0)
            m["wpb"].update(ntok)
            m["bsz"].update(nsent)
            m["gnorm"].update(gnorm)
            m["clip"].update(clipped)
            m["train_loss"].update(loss, sample_size)
            self._last_logging_output = {"loss": loss, "nll_loss": nll, "ntokens": ntok, "nsentences": nsent,
                                         "sample_size": sample_size, "gnorm": gnorm}

    def flush_stats(self):
        """Block until every posted step has reached the meters."""
        self._consume_stats(keep_latest=0)
        if self._optimizer is not None and hasattr(self._optimizer.optimizer, "mark_step") and self._arena is not None:
            self._optimizer.optimizer.mark_step()

    @torch.no_grad()
    def valid_step(self, sample, raise_oom=False):
        """Forward only; returns ``{'loss', 'sample_size', 'nsentences', 'ntokens'}`` summed over ranks."""
        model = self.model
        self.flush_pending_update()
        model.eval()
        sample = self._prepare_sample(sample)
        ignore = sample is None
        if ignore:
            sample = self._prepare_sample(self._dummy_batch)
        loss, sample_size, logging_output = self.task.valid_step(sample, model)
        vec = torch.zeros(4, dtype=torch.float64, device=self.device)
        if not ignore:
            vec[0] = _as_dev_scalar(logging_output.get("loss", 0.0), self.device) * sample_size
            vec[1:4] = torch.tensor([float(sample_size), float(logging_output.get("nsentences", 0)),
                                     float(logging_output.get("ntokens", 0))], dtype=torch.float64).to(self.device)
        if self.world_size > 1:
            torch.distributed.all_reduce(vec)
        tot, ss, ns, nt = vec.tolist()
        out = {"loss": tot / ss if ss > 0 else 0.0, "sample_size": ss, "nsentences": ns, "ntokens": nt}
        if ss > 0:
            self.meters["valid_loss"].update(out["loss"], ss)
        return out

    # ---- small accessors --------------------------------------------------------------------------------
    def zero_grad(self):
        self.optimizer.zero_grad()

    def lr_step(self, epoch, val_loss=None):
        self.lr_scheduler.step(epoch, val_loss)
        return self.lr_step_update()

    def lr_step_update(self):
        return self.lr_scheduler.step_update(self.get_num_updates())

    def get_lr(self):
        return self.optimizer.get_lr()

    def get_model(self):
        return self._model

    def get_meter(self, name):
        return self.meters.get(name)

    def get_num_updates(self):
        return self._num_updates

    def set_num_updates(self, num_updates):
        self._num_updates = num_updates
        self.lr_step_update()


def _as_dev_scalar(x, device) -> torch.Tensor:
    if torch.is_tensor(x):
        return x.detach().reshape(()).to(device=device, dtype=torch.float64)
    return torch.tensor(float(x), dtype=torch.float64, device=device)
