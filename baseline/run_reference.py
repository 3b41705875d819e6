"""Reference arm of bench.py: runs the UNMODIFIED reference installed in ``baseline/_ref`` (pip --no-deps
--target) through its own public API and stock code path —
``tasks.LanguageModelingTask`` -> ``Controller`` (DistributedDataParallel + NCCL) -> ``controller.train_step`` driven by
its own ``EpochBatchIterator`` / ``GroupedIterator`` exactly as its ``train()`` does — on the same synthetic corpus,
model config, batch size and optimizer flags as our arm.  None of this framework's models, kernels or engine
are on that path; the only additions are import stubs for packages the image lacks (``baseline/shims``: h5py
backed by .npz, boto3/botocore) and ``sys.path`` entries the reference needs (it imports its own modules by
bare name, e.g. hetseq/checkpoint_utils.py:15, hetseq/data/data_utils.py:50)."""
from __future__ import annotations

import json
import os
import shutil
import sys
import tempfile
import time
import warnings

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
REF = os.path.join(HERE, "_ref")


def _unavailable(why):
    print(json.dumps({"impl": "reference", "unavailable": why}), flush=True)
    return 0


_CONLL_TAGS = ["O", "B-PER", "I-PER", "B-ORG", "I-ORG", "B-LOC", "I-LOC", "B-MISC", "I-MISC"]


def install_conll_loader():
    """The reference's NER task calls ``datasets.load_dataset(<loading script>, data_files=...)``
    (hetseq/tasks/bert_for_token_classification_task.py:44).  The ``datasets`` release on this image no longer runs
    loading scripts, so — like the h5py stand-in — the harness supplies what the dependency cannot: a
    ``load_dataset`` that reads the CoNLL-2003 column files into the same DatasetDict (``tokens`` + ``ner_tags`` as
    a ClassLabel sequence).  Reference code is untouched; it keeps doing its own tokenisation, label alignment,
    collation and batching on the result."""
    import datasets

    def read(path):
        toks, tags, rows = [], [], {"id": [], "tokens": [], "ner_tags": []}
        with open(path) as f:
            for line in f:
                line = line.strip()
                if not line or line.startswith("-DOCSTART-"):
                    if toks:
                        rows["id"].append(str(len(rows["id"]))); rows["tokens"].append(toks); rows["ner_tags"].append(tags)
                        toks, tags = [], []
                    continue
                parts = line.split()
                toks.append(parts[0]); tags.append(_CONLL_TAGS.index(parts[-1]))
        if toks:
            rows["id"].append(str(len(rows["id"]))); rows["tokens"].append(toks); rows["ner_tags"].append(tags)
        return rows

    feats = datasets.Features({"id": datasets.Value("string"), "tokens": datasets.Sequence(datasets.Value("string")),
                               "ner_tags": datasets.Sequence(datasets.ClassLabel(names=_CONLL_TAGS))})

    def load_dataset(path=None, data_files=None, **kwargs):
        return datasets.DatasetDict({split: datasets.Dataset.from_dict(read(fn), features=feats) for split, fn in (data_files or {}).items()})

    datasets.load_dataset = load_dataset
    # datasets' torch formatter imports torchvision.io.VideoReader, which this image's torchvision no longer has
    datasets.config.TORCHVISION_AVAILABLE = False


def run_reference(a):
    if not os.path.isdir(os.path.join(REF, "hetseq")):
        return _unavailable("baseline/_ref is missing (pip install --no-deps --target baseline/_ref /root/reference)")
    try:
        import torch
    except Exception as e:
        return _unavailable("torch import failed: %s" % e)
    if not torch.cuda.is_available():
        return _unavailable("no CUDA device visible (the reference's Controller is CUDA-only)")
    warnings.filterwarnings("ignore")
    sys.path.insert(0, os.path.join(HERE, "shims"))
    sys.path.insert(0, REF)
    sys.path.insert(0, os.path.join(REF, "hetseq"))
    sys.path.insert(0, ROOT)
    import numpy as np
    import torch.distributed as dist
    from bench import ClockSampler, MODELS, PUBLISHED_SEQ_PER_SEC_PER_GPU, _metric_name, bench_config
    from hetseq_9cme_b200.data import synthetic   # data *generation* only (same arrays as our arm)

    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    torch.cuda.set_device(local)
    try:
        from hetseq import options, tasks, distributed_utils
        from hetseq.controller import Controller
        from hetseq.data import iterators
    except Exception as e:
        return _unavailable("reference import failed: %r" % (e,))

    tmp = tempfile.mkdtemp(prefix="refbench_r%d_" % rank)
    try:
        nsteps_needed = 2 * a.steps + a.warmup + 8       # fp32 pass + TF32 companion pass
        per_rank = nsteps_needed * a.batch
        max_pred = 20 if a.seq_len <= 128 else 80
        vocab_path = synthetic.write_vocab(os.path.join(tmp, "vocab.txt"))
        common = ["--max-sentences", str(a.batch), "--max-epoch", "1000", "--lr", "1e-4", "--warmup-updates", "100",
                  "--total-num-update", "100000", "--weight-decay", "0.01", "--clip-norm", "25", "--fast-stat-sync",
                  "--valid-subset", "test", "--save-dir", os.path.join(tmp, "ck"), "--no-save", "--log-format", "none",
                  "--distributed-world-size", str(world)]
        if getattr(a, "task", "bert") == "ner":
            # BASELINE config 5: token classification on CoNLL-2003-shaped text, the reference's own task / collator /
            # dataset wrapper (hetseq/tasks/bert_for_token_classification_task.py:189-228); same corpus as our arm
            install_conll_loader()
            import importlib
            paths = synthetic.make_conll(tmp, n_train=per_rank * world + a.batch, n_valid=a.batch, n_test=a.batch,
                                         vocab_words=20000, seed=1234)
            cfg_path = synthetic.write_bert_config(os.path.join(tmp, "bert_config.json"), vocab_size=20005, **MODELS[a.model])
            argv = ["--data", tmp, "--dict", paths["vocab"], "--config_file", cfg_path, "--train_file", paths["train"],
                    "--validation_file", paths["validation"], "--test_file", paths["test"], "--extension_file", "conll2003",
                    "--num-workers", "2", "--find-unused-parameters"] + common      # as its own script passes it (the pooler is unused)
            parser = options.get_training_parser(task="BertForTokenClassification", optimizer="adam",
                                                 lr_scheduler="PolynomialDecayScheduler")
            task_cls = importlib.import_module("hetseq.tasks.bert_for_token_classification_task").BertForTokenClassificationTask
            task_name = "BertForTokenClassification"
        else:
            n_per_file = (per_rank * world + 1) // 2 + a.batch
            for i in range(2):
                arrs = synthetic.bert_shard_arrays(n_per_file, a.seq_len, max_pred, 30522, 1234 + i)
                np.savez(os.path.join(tmp, "train_shard_%d.npz" % i), **arrs)
            np.savez(os.path.join(tmp, "test_shard_0.npz"), **synthetic.bert_shard_arrays(a.batch, a.seq_len, max_pred, 30522, 2234))
            cfg_path = synthetic.write_bert_config(os.path.join(tmp, "bert_config.json"), **MODELS[a.model])
            argv = ["--data", tmp, "--dict", vocab_path, "--config_file", cfg_path, "--num-workers", "4"] + common
            parser = options.get_training_parser(task="bert", optimizer="adam", lr_scheduler="PolynomialDecayScheduler")
            task_cls = tasks.LanguageModelingTask
            task_name = "bert"
        args = options.parse_args_and_arch(parser, argv)
        args.task = task_name
        args.device_id = local
        args.distributed_rank = rank
        np.random.seed(args.seed)
        torch.manual_seed(args.seed)
        if world > 1:
            args.distributed_init_method = "env://"
            args.distributed_rank = distributed_utils.distributed_init(args)
        task = task_cls.setup_task(args)
        if task_name == "bert":
            task.load_dataset("test", combine=False, epoch=0)
        model = task.build_model(args)
        controller = Controller(args, task, model)
        epoch_itr = controller.get_train_iterator(epoch=0, load_dataset=True)
        controller.lr_step(epoch_itr.epoch)
        itr = epoch_itr.next_epoch_itr(shuffle=True)
        itr = iterators.GroupedIterator(itr, 1)

        def sync_all():
            torch.cuda.synchronize()
            if world > 1:
                dist.barrier()
                torch.cuda.synchronize()

        def measure(nwarm):
            """K public-API steps (H2D copies and the reference's own host syncs included); device- and wall-timed."""
            for _ in range(nwarm):
                controller.train_step(next(itr))
            sync_all()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            h2d = 0
            t0 = time.perf_counter()
            e0.record()
            for _ in range(a.steps):
                samples = next(itr)
                first = samples[0]
                h2d += sum(t.numel() * t.element_size() for t in (first.values() if isinstance(first, dict) else first) if torch.is_tensor(t))
                controller.train_step(samples)
            e1.record()
            sync_all()
            wall_ms = (time.perf_counter() - t0) * 1e3
            t = torch.tensor([e0.elapsed_time(e1), wall_ms], device="cuda", dtype=torch.float64)
            if world > 1:
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
            return t.tolist() + [h2d]

        # headline: the reference exactly as it ships (torch default: fp32 matmuls on the FP32 pipe)
        sampler = ClockSampler(local)
        sampler.start()
        dev_ms, wall_ms, h2d = measure(a.warmup)
        clocks = sampler.stop()
        loss_meter = controller.get_meter("train_loss")
        final_loss = float(loss_meter.avg) if loss_meter is not None else None
        # companion datum: same code, cuBLAS allowed to use TF32 tensor cores (a torch switch in this harness, nothing
        # in baseline/_ref is touched) — the like-for-like partner of our TF32 arm
        tf32 = None
        if getattr(a, "precision", "tf32") in ("tf32", "bf16") and True:
            torch.backends.cuda.matmul.allow_tf32 = True
            torch.backends.cudnn.allow_tf32 = True
            try:
                d2, w2, _ = measure(3)
                tf32 = {"value": round(a.batch * world * a.steps / (d2 / 1e3), 2), "ms_per_step": round(d2 / a.steps, 4),
                        "e2e_value": round(a.batch * world * a.steps / (w2 / 1e3), 2), "allow_tf32": True}
            except StopIteration:
                tf32 = None
            torch.backends.cuda.matmul.allow_tf32 = False
            torch.backends.cudnn.allow_tf32 = False
        gb = a.batch * world
        value = gb * a.steps / (dev_ms / 1e3)
        out = {"metric": _metric_name(a.model, a.seq_len, getattr(a, "task", "bert")), "value": round(value, 2), "unit": "sequences/s",
               "n_gpus": world, "steps": a.steps, "warmup": a.warmup, "ms_per_step": round(dev_ms / a.steps, 4),
               "higher_is_better": True, "scaling": "weak", "vs_baseline": round(value / (PUBLISHED_SEQ_PER_SEC_PER_GPU * world), 2),
               "dtype": "fp32", "data": "synthetic", "impl": "reference",
               "config": bench_config(a, world, max_pred),
               "impl_detail": {"allow_tf32": False,
                               "path": "hetseq.controller.Controller.train_step (DDP + NCCL + cuBLAS), unmodified",
                               "timing": "CUDA events around the K public-API steps incl. their H2D copies and host syncs; max over ranks"},
               "reference_tf32": tf32,
               "clocks": clocks, "final_loss": (round(final_loss, 4) if final_loss is not None else None),
               "e2e": {"value": round(gb * a.steps / (wall_ms / 1e3), 2), "unit": "sequences/s",
                       "ms_per_step": round(wall_ms / a.steps, 4), "h2d_bytes_per_step": int(h2d / a.steps),
                       "d2h_bytes_per_step": 48},
               "gpu_launches": 0}
        if rank == 0 or world == 1:
            sys.__stdout__.write(json.dumps(out) + "\n")
            sys.__stdout__.flush()
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
    except Exception as e:
        import traceback
        traceback.print_exc(file=sys.stderr)
        return _unavailable("reference run failed: %r" % (e,))
    finally:
        shutil.rmtree(tmp, ignore_errors=True)
    return 0
