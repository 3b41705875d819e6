#!/usr/bin/env python3
"""Command-line entry point: ``python hetseq_9cme_b200/train.py --task ... `` (one command per node).

Same launch model and flags as hetseq/train.py:196-250 — (a) ``--distributed-init-method`` given: spawn
``--distributed-gpus`` processes on this node with ranks ``--distributed-rank + i``; (b) no init method but world
size > 1: single node, rendezvous on a free local port; (c) single process — and the same epoch loop
(hetseq/train.py:25-168), with the parts the reference left unreachable now working: validation,
``--save-interval-updates`` mid-epoch checkpoints, ``checkpoint_best.pt``, and resume from ``checkpoint_last.pt``.
"""
from __future__ import annotations

import collections
import math
import os
import sys

# allow `python hetseq_9cme_b200/train.py` — also in the children torch.multiprocessing.spawn starts, which import this
# file again as "__mp_main__"
if __name__ in ("__main__", "__mp_main__") and not __package__:
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    __package__ = "hetseq_9cme_b200"

import numpy as np
import torch

from . import checkpoint_utils, distributed_utils, options, progress_bar, tasks, utils
from .controller import Controller
from .data import iterators
from .meters import AverageMeter, StopwatchMeter


def _maybe_fault_inject(args, num_updates):
    spec = getattr(args, "fault_inject", None) or os.environ.get("HETSEQ_FAULT_INJECT")
    if spec:
        rank, step = (int(x) for x in spec.split(":"))
        # only in the first attempt: a restarted run passes the same update count again
        if rank == getattr(args, "distributed_rank", 0) and num_updates == step and not getattr(args, "restart_attempt", 0):
            distributed_utils.log_all("| fault injection: rank {} exits at update {}".format(rank, step))
            os._exit(17)


def main(args, init_distributed=False):
    assert args.max_tokens is not None or args.max_sentences is not None, \
        "Must specify batch size either with --max-tokens or --max-sentences"
    use_cuda = torch.cuda.is_available() and not args.cpu
    if use_cuda:
        torch.cuda.set_device(args.device_id)
    np.random.seed(args.seed)
    torch.manual_seed(args.seed)
    if init_distributed:
        args.distributed_rank = distributed_utils.distributed_init(args)
    if distributed_utils.is_master(args):
        checkpoint_utils.verify_checkpoint_directory(args.save_dir)
    print(args, flush=True)

    if args.task not in tasks.TASK_REGISTRY:
        raise ValueError("unknown task {}".format(args.task))
    task = tasks.TASK_REGISTRY[args.task].setup_task(args)
    for valid_sub_split in args.valid_subset.split(","):
        task.load_dataset(valid_sub_split, combine=False, epoch=0)
    model = task.build_model(args)
    print("| num. model params: {} (num. trained: {})".format(
        sum(p.numel() for p in model.parameters()), sum(p.numel() for p in model.parameters() if p.requires_grad)))
    controller = Controller(args, task, model)
    print("| training on {} {}".format(args.distributed_world_size, "GPUs" if use_cuda else "CPU processes"))
    print("| max tokens per GPU = {} and max sentences per GPU = {}".format(args.max_tokens, args.max_sentences))

    if hasattr(checkpoint_utils.save_checkpoint, "best"):      # per-run state (matters when main() is called twice)
        del checkpoint_utils.save_checkpoint.best
    extra_state, epoch_itr = checkpoint_utils.load_checkpoint(args, controller)

    max_epoch = args.max_epoch or math.inf
    max_update = args.max_update or math.inf
    lr = controller.get_lr()
    train_meter = StopwatchMeter()
    train_meter.start()
    valid_subsets = args.valid_subset.split(",")
    while (lr > args.min_lr
           and (epoch_itr.epoch < max_epoch or (epoch_itr.epoch == max_epoch and epoch_itr._next_epoch_itr is not None))
           and controller.get_num_updates() < max_update):
        train(args, controller, task, epoch_itr)
        if not args.disable_validation and epoch_itr.epoch % args.validate_interval == 0:
            valid_losses = validate(args, controller, task, epoch_itr, valid_subsets)
        else:
            valid_losses = [None]
        lr = controller.lr_step(epoch_itr.epoch, valid_losses[0])
        if epoch_itr.epoch % args.save_interval == 0:
            checkpoint_utils.save_checkpoint(args, controller, epoch_itr, valid_losses[0])
        reload_dataset = ":" in (getattr(args, "data", "") or "")
        epoch_itr = controller.get_train_iterator(epoch_itr.epoch, load_dataset=reload_dataset)
    train_meter.stop()
    controller.flush_stats()
    print("| done training in {:.1f} seconds".format(train_meter.sum))
    return controller


def train(args, controller, task, epoch_itr):
    """One epoch."""
    update_freq = args.update_freq[epoch_itr.epoch] if epoch_itr.epoch < len(args.update_freq) else args.update_freq[-1]
    itr = epoch_itr.next_epoch_itr(fix_batches_to_gpus=args.fix_batches_to_gpus, shuffle=(epoch_itr.epoch >= args.curriculum))
    itr = iterators.GroupedIterator(itr, update_freq)
    progress = progress_bar.build_progress_bar(args, itr, epoch_itr.epoch, no_progress_bar="simple")
    extra_meters = collections.defaultdict(AverageMeter)
    valid_subsets = args.valid_subset.split(",")
    max_update = args.max_update or math.inf
    for i, samples in enumerate(progress, start=epoch_itr.iterations_in_epoch):
        try:
            log_output = controller.train_step(samples)
        except RuntimeError:
            # a peer that died (or stalled past B200_P2P_TIMEOUT_S) trips the reduction kernels' watchdog: say which
            # rank went missing and leave with a distinct exit code instead of hanging in the next collective
            report = controller.watchdog_report()
            if report:
                distributed_utils.log_all("| FATAL (rank {}): {}".format(getattr(args, "distributed_rank", 0), report))
                sys.stdout.flush()
                os._exit(86)
            raise
        if log_output is None:
            continue
        stats = get_training_stats(controller)
        for k, v in log_output.items():
            if k in ("loss", "nll_loss", "ntokens", "nsentences", "sample_size", "gnorm"):
                continue
            if "loss" in k or k == "accuracy":
                extra_meters[k].update(v, log_output["sample_size"])
            else:
                extra_meters[k].update(v)
            stats[k] = extra_meters[k].avg
        progress.log(stats, tag="train", step=stats["num_updates"])
        if i == 0:   # the first step pays for lazy initialisation; keep it out of wps / ups
            controller.get_meter("wps").reset()
            controller.get_meter("ups").reset()
        num_updates = controller.get_num_updates()
        _maybe_fault_inject(args, num_updates)
        if (not args.disable_validation and args.save_interval_updates > 0 and num_updates % args.save_interval_updates == 0
                and num_updates > 0):
            valid_losses = validate(args, controller, task, epoch_itr, valid_subsets)
            checkpoint_utils.save_checkpoint(args, controller, epoch_itr, valid_losses[0])
        elif args.save_interval_updates > 0 and num_updates % args.save_interval_updates == 0 and num_updates > 0:
            checkpoint_utils.save_checkpoint(args, controller, epoch_itr, None)
        if num_updates >= max_update:
            break
    controller.flush_stats()
    stats = get_training_stats(controller)
    for k, meter in extra_meters.items():
        stats[k] = meter.avg
    progress.print(stats, tag="train", step=stats["num_updates"])
    for k in ("train_loss", "train_nll_loss", "wps", "ups", "wpb", "bsz", "gnorm", "clip"):
        meter = controller.get_meter(k)
        if meter is not None:
            meter.reset()


def validate(args, controller, task, epoch_itr, subsets):
    """Evaluate on each validation subset; returns the list of losses."""
    losses = []
    for subset in subsets:
        try:
            dataset = task.dataset(subset)
        except KeyError:
            losses.append(None)
            continue
        itr = task.get_batch_iterator(
            dataset=dataset, max_tokens=args.max_tokens_valid, max_sentences=args.max_sentences_valid,
            required_batch_size_multiple=args.required_batch_size_multiple, seed=args.seed,
            num_shards=args.distributed_world_size, shard_id=getattr(args, "distributed_rank", 0) or 0,
            num_workers=args.num_workers).next_epoch_itr(shuffle=False)
        meter = controller.get_meter("valid_loss")
        meter.reset()
        controller._dummy_batch = controller._dummy_batch or None
        for sample in itr:
            if controller._dummy_batch is None and sample is not None:
                controller._dummy_batch = sample
            controller.valid_step(sample)
        stats = collections.OrderedDict()
        stats["loss"] = meter.avg
        stats["ppl"] = utils.get_perplexity(meter.avg)
        stats["num_updates"] = controller.get_num_updates()
        print("| epoch {:03d} | valid on '{}' subset | loss {:.3f} | ppl {} | num_updates {}".format(
            epoch_itr.epoch, subset, meter.avg, stats["ppl"], stats["num_updates"]), flush=True)
        if getattr(args, "log_jsonl", None) and distributed_utils.is_master(args):
            progress_bar.json_progress_bar([], epoch_itr.epoch, path=args.log_jsonl).print(stats, tag="valid/" + subset)
        losses.append(meter.avg if meter.count > 0 else None)
    return losses or [None]


def get_training_stats(controller):
    stats = collections.OrderedDict()
    stats["loss"] = controller.get_meter("train_loss")
    if controller.get_meter("train_nll_loss").count > 0:
        nll_loss = controller.get_meter("train_nll_loss")
        stats["nll_loss"] = nll_loss
    else:
        nll_loss = controller.get_meter("train_loss")
    stats["ppl"] = utils.get_perplexity(nll_loss.avg)
    stats["wps"] = controller.get_meter("wps")
    stats["ups"] = controller.get_meter("ups")
    stats["wpb"] = controller.get_meter("wpb")
    stats["bsz"] = controller.get_meter("bsz")
    stats["num_updates"] = controller.get_num_updates()
    stats["lr"] = controller.get_lr()
    stats["gnorm"] = controller.get_meter("gnorm")
    stats["clip"] = controller.get_meter("clip")
    stats["oom"] = controller.get_meter("oom") if controller.get_meter("oom") is not None else 0
    if controller.get_meter("loss_scale") is not None:
        stats["loss_scale"] = controller.get_meter("loss_scale")
    stats["wall"] = round(controller.get_meter("wall").elapsed_time)
    stats["train_wall"] = controller.get_meter("train_wall")
    return stats


def distributed_main(i, args, start_rank=0):
    args.device_id = i
    if args.distributed_rank is None:
        args.distributed_rank = start_rank + i
    main(args, init_distributed=True)


def _spawn_supervised(args, nprocs, spawn_args, fresh_port=False):
    """``torch.multiprocessing.spawn`` with failure recovery (``--max-restarts``): when a rank dies, spawn stops its
    siblings and raises; all ranks of this node are then started again and resume from ``checkpoint_last.pt`` (the
    checkpoint carries the iterator position, the schedule and the update count, so the run continues where the last
    checkpoint left it).  The reference has no recovery path: a dead rank hangs the job until the scheduler kills it.
    On several nodes every node's launcher sees its own ranks fail (collective error, or the peer-to-peer watchdog,
    exit code 86) and restarts them towards the same ``--distributed-init-method`` (only the single-node case is
    covered by a test)."""
    from torch.multiprocessing import ProcessExitedException, ProcessRaisedException
    attempt = 0
    while True:
        args.restart_attempt = attempt
        if fresh_port:
            args.distributed_init_method = "tcp://127.0.0.1:{}".format(distributed_utils.find_free_port())
        try:
            torch.multiprocessing.spawn(fn=distributed_main, args=spawn_args, nprocs=nprocs)
            return
        except (ProcessExitedException, ProcessRaisedException) as e:
            if attempt >= int(getattr(args, "max_restarts", 0) or 0):
                raise
            attempt += 1
            first = str(e).strip().splitlines()[0] if str(e).strip() else type(e).__name__
            print("| a rank failed ({}); restart {}/{} from the last checkpoint".format(first, attempt, args.max_restarts),
                  flush=True)


def cli_main(argv=None):
    args = options.parse_training_args(argv)
    ndev = torch.cuda.device_count() if (torch.cuda.is_available() and not args.cpu) else 0
    if args.distributed_init_method is not None:
        if ndev:
            assert args.distributed_gpus <= ndev, "--distributed-gpus exceeds the visible GPUs"
        if args.distributed_gpus > 1 and not args.distributed_no_spawn:
            start_rank = args.distributed_rank
            args.distributed_rank = None
            _spawn_supervised(args, args.distributed_gpus, (args, start_rank))
        else:
            distributed_main(args.device_id, args)
    elif args.distributed_world_size > 1:
        if ndev:
            assert args.distributed_world_size <= ndev, "--distributed-world-size exceeds the visible GPUs"
        args.distributed_rank = None
        _spawn_supervised(args, args.distributed_world_size, (args,), fresh_port=True)
    else:
        main(args)


if __name__ == "__main__":
    cli_main()
