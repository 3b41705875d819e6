#!/usr/bin/env python3
"""Loss-curve parity against the reference on one GPU (SURVEY §7.6 "Integration").

    python tools/loss_parity.py all OUT [--steps 20] [--layers 12] [--precision tf32|bf16]

1. ``ref``  (own process): the UNMODIFIED reference from baseline/_ref builds BERT-base with its own initialisation,
   writes a checkpoint with ITS save code, then runs K training steps (fp32, dropout 0, its Controller / Adam /
   scheduler) and records per-step loss and gradient norm.
2. ``ours`` (own process): this framework restores the reference's checkpoint with its ordinary resume path (checkpoint
   compatibility is part of the contract), runs the same K steps on the same batches through the native kernels
   (CUDA graph and packed batches on, as a user gets them) and records the same two series.
3. ``all``: runs both and compares step by step.

Data: the same synthetic shard arrays, as ``.npz`` for the reference (through its h5py stand-in) and ``.b2s`` for us.
"""
import argparse
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

BASE = dict(hidden_size=768, num_attention_heads=12, intermediate_size=3072, hidden_dropout_prob=0.0,
            attention_probs_dropout_prob=0.0)
BATCH, SEQ, PRED, VOCAB = 32, 128, 20, 30522
COMMON = ["--max-sentences", str(BATCH), "--max-epoch", "1000", "--lr", "1e-4", "--warmup-updates", "5", "--total-num-update", "1000",
          "--weight-decay", "0.01", "--clip-norm", "1.0", "--fast-stat-sync", "--valid-subset", "test", "--num-workers", "0",
          "--log-format", "none", "--distributed-world-size", "1", "--seed", "11"]


def make_data(out, steps):
    import numpy as np
    from hetseq_9cme_b200.data import shard_io, synthetic
    n = (steps + 4) * BATCH // 2
    for d in ("ref", "ours"):
        os.makedirs(os.path.join(out, d), exist_ok=True)
    for name, count, seed in (("train_shard_0", n, 1), ("train_shard_1", n, 2), ("test_shard_0", BATCH, 3)):
        arrs = synthetic.bert_shard_arrays(count, SEQ, PRED, VOCAB, seed)
        np.savez(os.path.join(out, "ref", name + ".npz"), **arrs)
        shard_io.write_b2shard(os.path.join(out, "ours", name + ".b2s"), arrs)
    synthetic.write_vocab(os.path.join(out, "vocab.txt"))


def run_ref(out, steps, layers):
    import warnings
    warnings.filterwarnings("ignore")
    base = os.path.join(ROOT, "baseline")
    ref = os.path.join(base, "_ref")
    for p in (os.path.join(base, "shims"), ref, os.path.join(ref, "hetseq")):
        sys.path.insert(0, p)
    import numpy as np
    import torch
    from hetseq_9cme_b200.data import synthetic
    from hetseq import options, tasks
    from hetseq.controller import Controller
    from hetseq.data import iterators
    cfg = synthetic.write_bert_config(os.path.join(out, "bert_config.json"), num_hidden_layers=layers, **BASE)
    argv = ["--data", os.path.join(out, "ref"), "--dict", os.path.join(out, "vocab.txt"), "--config_file", cfg,
            "--save-dir", os.path.join(out, "ck_ref")] + COMMON
    parser = options.get_training_parser(task="bert", optimizer="adam", lr_scheduler="PolynomialDecayScheduler")
    args = options.parse_args_and_arch(parser, argv)
    args.device_id, args.distributed_rank = 0, 0
    torch.cuda.set_device(0)
    np.random.seed(args.seed); torch.manual_seed(args.seed)
    task = tasks.LanguageModelingTask.setup_task(args)
    task.load_dataset("test", combine=False, epoch=0)
    model = task.build_model(args)
    c = Controller(args, task, model)
    it = c.get_train_iterator(epoch=0, load_dataset=True)
    c.lr_step(it.epoch)
    os.makedirs(args.save_dir, exist_ok=True)
    c.save_checkpoint(os.path.join(out, "ref_init.pt"), {"train_iterator": {"epoch": 0, "iterations_in_epoch": 0}, "val_loss": None})
    itr = iterators.GroupedIterator(it.next_epoch_itr(shuffle=False), 1)
    series = []
    for _ in range(steps):
        lo = c.train_step(next(itr))
        series.append({"loss": float(lo["loss"]), "gnorm": float(c.get_meter("gnorm").val)})
    json.dump(series, open(os.path.join(out, "ref_series.json"), "w"))
    print("reference:", " ".join("%.4f" % s["loss"] for s in series))


def run_ours(out, steps, layers, precision):
    import torch
    from hetseq_9cme_b200 import checkpoint_utils, options, tasks
    from hetseq_9cme_b200.controller import Controller
    from hetseq_9cme_b200.data import synthetic
    cfg = synthetic.write_bert_config(os.path.join(out, "bert_config.json"), num_hidden_layers=layers, **BASE)
    argv = ["--task", "bert", "--data", os.path.join(out, "ours"), "--dict", os.path.join(out, "vocab.txt"), "--config_file", cfg,
            "--save-dir", os.path.join(out, "ck_ours"), "--restore-file", os.path.join(out, "ref_init.pt"), "--reset-optimizer",
            "--reset-dataloader", "--reset-meters", "--precision", precision, "--sync-stats",
            "--max-predictions-per-seq", str(PRED)] + COMMON
    args = options.parse_training_args(argv)
    torch.cuda.set_device(0)
    torch.manual_seed(args.seed)
    task = tasks.LanguageModelingTask.setup_task(args)
    task.load_dataset("test")
    c = Controller(args, task, task.build_model(args))
    extra, it = checkpoint_utils.load_checkpoint(args, c)
    assert extra is not None, "the reference's checkpoint did not load"
    itr = it.next_epoch_itr(shuffle=False)
    series = []
    for _ in range(steps):
        lo = c.train_step([next(itr)])
        series.append({"loss": float(lo["loss"]), "gnorm": float(lo.get("gnorm", c.get_meter("gnorm").val))})
    c.flush_stats()
    json.dump(series, open(os.path.join(out, "ours_series_%s.json" % precision), "w"))
    print("ours (%s, graph %s):" % (precision, c._graph is not None), " ".join("%.4f" % s["loss"] for s in series))


def compare(out, precision, tol_first, tol_last):
    ref = json.load(open(os.path.join(out, "ref_series.json")))
    ours = json.load(open(os.path.join(out, "ours_series_%s.json" % precision)))
    worst = 0.0
    print("step   ref loss  our loss   rel      ref gnorm  our gnorm")
    for i, (r, o) in enumerate(zip(ref, ours)):
        rel = abs(o["loss"] - r["loss"]) / abs(r["loss"])
        worst = max(worst, rel)
        print("%3d  %9.5f %9.5f  %.2e  %9.4f  %9.4f" % (i, r["loss"], o["loss"], rel, r["gnorm"], o["gnorm"]))
    first = abs(ours[0]["loss"] - ref[0]["loss"]) / abs(ref[0]["loss"])
    g0 = abs(ours[0]["gnorm"] - ref[0]["gnorm"]) / abs(ref[0]["gnorm"])
    ok = first < tol_first and worst < tol_last and g0 < 5 * tol_first and ref[-1]["loss"] < ref[0]["loss"]
    print("LOSS PARITY %s (%s): first-step rel %.2e, worst rel %.2e over %d steps, first gnorm rel %.2e" % (
        "PASS" if ok else "FAIL", precision, first, worst, len(ref), g0))
    return ok


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["all", "ref", "ours", "compare"])
    ap.add_argument("out")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--layers", type=int, default=12)
    ap.add_argument("--precision", default="tf32")
    a = ap.parse_args()
    os.makedirs(a.out, exist_ok=True)
    tol = {"tf32": (2e-3, 1e-2), "bf16": (1e-2, 4e-2)}[a.precision]
    if a.what == "ref":
        return run_ref(a.out, a.steps, a.layers)
    if a.what == "ours":
        return run_ours(a.out, a.steps, a.layers, a.precision)
    if a.what == "compare":
        return 0 if compare(a.out, a.precision, *tol) else 1
    make_data(a.out, a.steps)
    for what in ("ref", "ours"):
        r = subprocess.run([sys.executable, os.path.abspath(__file__), what, a.out, "--steps", str(a.steps), "--layers", str(a.layers),
                            "--precision", a.precision], cwd=ROOT)
        if r.returncode != 0:
            print("LOSS PARITY FAIL: the %s arm exited with %d" % (what, r.returncode))
            return 1
    return 0 if compare(a.out, a.precision, *tol) else 1


if __name__ == "__main__":
    sys.exit(main())
