"""Checkpoint files: names, schema, save/load, pruning (API of hetseq/checkpoint_utils.py:14-222).

On-disk compatibility with the reference is kept exactly — file names ``checkpoint{epoch}.pt``,
``checkpoint_{epoch}_{updates}.pt``, ``checkpoint_best.pt``, ``checkpoint_last.pt``; one ``torch.save``d dict with
keys ``args / model / optimizer_history / extra_state / last_optimizer_state``; model keys without any wrapper
prefix; ``optimizer_name`` = wrapper class name (``_Adam``) — so files move between the two code bases.

Fixed relative to the reference: ``extra_state`` is actually written (the reference saves ``{}`` and
can therefore never resume), loading tolerates an empty one, writes are atomic (tmp + rename) and retried, all
ranks meet at a barrier after the master has written, un-picklable / huge task objects are stripped from
``args``, and mid-epoch ``--save-interval-updates`` checkpoints are reachable.
"""
from __future__ import annotations

import argparse
import collections
import os
import re
import shutil
import time
import traceback
from typing import Any, Dict, List, Optional

import torch

from . import distributed_utils, meters as _meters

_ARG_DROP = ("tokenized_datasets", "tokenizer", "data_collator", "label_list", "EntNameID")


def clean_args(args) -> argparse.Namespace:
    """Copy of ``args`` holding only plain picklable values."""
    out = argparse.Namespace()
    for k, v in vars(args).items():
        if k in _ARG_DROP:
            continue
        if isinstance(v, (int, float, str, bool, type(None), list, tuple, dict)):
            setattr(out, k, v)
    return out


def checkpoint_paths(path: str, pattern: str = r"checkpoint(\d+)\.pt") -> List[str]:
    """Files in ``path`` matching ``pattern``, sorted by the first group, descending."""
    pt = re.compile(pattern)
    entries = []
    for i, f in enumerate(os.listdir(path)):
        m = pt.fullmatch(f)
        if m is not None:
            idx = int(m.group(1)) if len(m.groups()) > 0 else i
            entries.append((idx, m.group(0)))
    return [os.path.join(path, x[1]) for x in sorted(entries, reverse=True)]


def torch_persistent_save(obj, filename: str, attempts: int = 3) -> None:
    """Atomic ``torch.save`` with retries (shared file systems fail transiently)."""
    tmp = filename + ".tmp.%d" % os.getpid()
    for i in range(attempts):
        try:
            torch.save(obj, tmp)
            os.replace(tmp, filename)
            return
        except Exception:
            if i == attempts - 1:
                print(traceback.format_exc())
                raise
            time.sleep(0.2 * (i + 1))


def convert_state_dict_type(state_dict, ttype=torch.FloatTensor):
    """Recursively cast tensors (optimizer state is stored as CPU fp32)."""
    if isinstance(state_dict, dict):
        return collections.OrderedDict((k, convert_state_dict_type(v, ttype)) for k, v in state_dict.items())
    if isinstance(state_dict, list):
        return [convert_state_dict_type(v, ttype) for v in state_dict]
    if torch.is_tensor(state_dict):
        return state_dict.detach().to("cpu").type(ttype) if state_dict.is_floating_point() else state_dict.detach().cpu()
    return state_dict


def save_state(filename, args, model_state_dict, criterion, optimizer, lr_scheduler, num_updates, optim_history=None,
               extra_state=None):
    if optim_history is None:
        optim_history = []
    if extra_state is None:
        extra_state = {}
    state = {
        "args": clean_args(args),
        "model": convert_state_dict_type(model_state_dict) if model_state_dict else {},
        "optimizer_history": optim_history + [{
            "optimizer_name": optimizer.__class__.__name__,
            "lr_scheduler_state": lr_scheduler.state_dict(),
            "num_updates": num_updates,
        }],
        "extra_state": extra_state,
    }
    if not getattr(args, "no_save_optimizer_state", False):
        state["last_optimizer_state"] = convert_state_dict_type(optimizer.state_dict())
    torch_persistent_save(state, filename)


def load_checkpoint_to_cpu(path: str) -> Dict[str, Any]:
    # the pickled argparse.Namespace needs weights_only=False on torch >= 2.6
    return torch.load(path, map_location="cpu", weights_only=False)


def verify_checkpoint_directory(save_dir: str) -> None:
    os.makedirs(save_dir, exist_ok=True)
    probe = os.path.join(save_dir, "dummy")
    try:
        with open(probe, "w"):
            pass
    except OSError as e:
        print("| Unable to access checkpoint save directory: {}".format(save_dir))
        raise e
    else:
        os.remove(probe)


def save_checkpoint(args, controller, epoch_itr, val_loss):
    """Write the checkpoints that are due now (epoch end or ``--save-interval-updates``), prune old ones."""
    if getattr(args, "no_save", False):
        return
    master = distributed_utils.is_master(args)
    if hasattr(controller, "prepare_checkpoint"):
        controller.prepare_checkpoint()       # every rank: sharded optimizer state is gathered before the master writes
    prev_best = getattr(save_checkpoint, "best", val_loss)
    if val_loss is not None:
        better = max if getattr(args, "maximize_best_checkpoint_metric", False) else min
        save_checkpoint.best = better(val_loss, prev_best) if prev_best is not None else val_loss

    if master:
        epoch = epoch_itr.epoch
        end_of_epoch = epoch_itr.end_of_epoch()
        updates = controller.get_num_updates()
        conds = collections.OrderedDict()
        conds["checkpoint{}.pt".format(epoch)] = (end_of_epoch and not args.no_epoch_checkpoints
                                                  and epoch % args.save_interval == 0)
        conds["checkpoint_{}_{}.pt".format(epoch, updates)] = (not end_of_epoch and args.save_interval_updates > 0
                                                               and updates % args.save_interval_updates == 0)
        conds["checkpoint_best.pt"] = (val_loss is not None and (not hasattr(save_checkpoint, "best")
                                                                 or save_checkpoint.best == val_loss))
        conds["checkpoint_last.pt"] = not args.no_last_checkpoints
        extra_state = {"train_iterator": epoch_itr.state_dict(), "val_loss": val_loss}
        if hasattr(save_checkpoint, "best"):
            extra_state["best"] = save_checkpoint.best
        names = [os.path.join(args.save_dir, fn) for fn, cond in conds.items() if cond]
        if names:
            t0 = time.perf_counter()
            controller.save_checkpoint(names[0], extra_state)
            for other in names[1:]:
                shutil.copyfile(names[0], other)
            print("| saved checkpoint {} (epoch {} @ {} updates) (writing took {:.2f} seconds)".format(
                names[0], epoch, updates, time.perf_counter() - t0))
        if not end_of_epoch and args.keep_interval_updates > 0:
            for old in checkpoint_paths(args.save_dir, r"checkpoint_\d+_(\d+)\.pt")[args.keep_interval_updates:]:
                if os.path.lexists(old):
                    os.remove(old)
        if args.keep_last_epochs > 0:
            for old in checkpoint_paths(args.save_dir, r"checkpoint(\d+)\.pt")[args.keep_last_epochs:]:
                if os.path.lexists(old):
                    os.remove(old)
    # nobody races into the next epoch while the file is half written (the reference has no barrier here)
    distributed_utils.barrier()


def load_checkpoint(args, controller):
    """Restore from ``--restore-file`` if present; returns ``(extra_state, epoch_itr)``."""
    if getattr(args, "distributed_rank", 0) in (0, None):
        os.makedirs(args.save_dir, exist_ok=True)
    path = args.restore_file if os.path.isabs(args.restore_file) else os.path.join(args.save_dir, args.restore_file)
    overrides = args.optimizer_overrides
    if isinstance(overrides, str):
        from .options import _literal
        overrides = _literal(overrides)
    extra_state = controller.load_checkpoint(path, args.reset_optimizer, args.reset_lr_scheduler, overrides,
                                             reset_meters=args.reset_meters)
    if extra_state is not None and "best" in extra_state and not args.reset_optimizer and not args.reset_meters:
        save_checkpoint.best = extra_state["best"]
    itr_state = (extra_state or {}).get("train_iterator")
    if itr_state is not None and not args.reset_dataloader:
        epoch_itr = controller.get_train_iterator(epoch=itr_state["epoch"], load_dataset=True)
        epoch_itr.load_state_dict(itr_state)
    else:
        epoch_itr = controller.get_train_iterator(epoch=0, load_dataset=True)
    controller.lr_step(epoch_itr.epoch)
    return extra_state, epoch_itr
