"""Command-line flag system.

Same two-stage design and the same flag names/defaults as the reference (``hetseq/options.py:5-383``,
flag table in SURVEY §5.6), so every shipped launch script parses unchanged.  The implementation is
table driven: each group is a list of ``(flags, kwargs)`` rows, task / optimizer / scheduler specific rows
are selected by name.  Flags that exist only in this framework are grouped under "B200 engine".
"""
from __future__ import annotations

import argparse
import ast
from typing import Callable, Iterable, List, Optional, Sequence

TASKS = ("bert", "mnist", "BertForTokenClassification", "BertForELClassification")
OPTIMIZERS = ("adam", "adadelta")
LR_SCHEDULERS = ("PolynomialDecayScheduler",)


# -------------------------------------------------------------------------------------------------
# small parsers
# -------------------------------------------------------------------------------------------------
def _literal(text):
    """Evaluate a Python literal (tuple/list/number/dict); falls back to ``eval`` for expressions such
    as ``1e-4*2`` that shipped scripts occasionally use."""
    try:
        return ast.literal_eval(text)
    except (ValueError, SyntaxError):
        return eval(text)  # noqa: S307 - parity with the reference, which evals these flags


def eval_str_list(x, type: Callable = float) -> Optional[List]:
    """``"0.1,0.2"`` / ``"[1,2]"`` / ``3`` -> list of ``type`` (ref: hetseq/options.py:355-363)."""
    if x is None:
        return None
    if isinstance(x, str):
        x = _literal(x)
    try:
        return [type(v) for v in x]
    except TypeError:
        return [type(x)]


def eval_bool(x, default: bool = False) -> bool:
    """Parse a textual boolean (ref: hetseq/options.py:366-372)."""
    if x is None:
        return default
    if isinstance(x, bool):
        return x
    try:
        return bool(_literal(x))
    except (TypeError, NameError, SyntaxError):
        return default


def _default_world_size() -> int:
    try:
        import torch
        return max(1, torch.cuda.device_count())
    except Exception:  # pragma: no cover - torch always importable here
        return 1


def _add_rows(target, rows: Iterable):
    for flags, kw in rows:
        target.add_argument(*flags, **kw)


# -------------------------------------------------------------------------------------------------
# flag tables
# -------------------------------------------------------------------------------------------------
_GENERAL = [
    (("--no-progress-bar",), dict(action="store_true", help="disable progress bar")),
    (("--seed",), dict(default=19940802, type=int, metavar="N", help="pseudo random number generator seed")),
    (("--cpu",), dict(action="store_true", help="use CPU instead of CUDA")),
    (("--log-interval",), dict(type=int, default=1, metavar="N", help="log progress every N batches")),
    (("--log-format",), dict(default="simple", choices=["none", "simple"], help="log format to use")),
]

_DATA_COMMON = [
    (("--num-workers",), dict(default=-1, type=int, metavar="N",
                              help="data-loading worker threads (<=0: load in the training process)")),
    (("--max-tokens",), dict(type=int, metavar="N", help="maximum number of tokens in a batch")),
    (("--max-sentences", "--batch-size"), dict(type=int, metavar="N", help="maximum number of sentences in a batch")),
    (("--required-batch-size-multiple",), dict(default=1, type=int, metavar="N",
                                               help="batch size will be a multiplier of this value")),
]

_DATA_TRAIN = [
    (("--train-subset",), dict(default="train", metavar="SPLIT", choices=["train", "valid", "test"],
                               help="data subset to use for training")),
    (("--valid-subset",), dict(default="valid", metavar="SPLIT",
                               help="comma separated list of data subsets to use for validation")),
    (("--validate-interval",), dict(type=int, default=1, metavar="N", help="validate every N epochs")),
    (("--disable-validation",), dict(action="store_true", help="disable validation")),
    (("--max-tokens-valid",), dict(type=int, metavar="N", help="max tokens in a validation batch")),
    (("--max-sentences-valid",), dict(type=int, metavar="N", help="max sentences in a validation batch")),
    (("--curriculum",), dict(default=0, type=int, metavar="N", help="don't shuffle batches for first N epochs")),
]

_BERT_FILES = [
    (("--dict",), dict(type=str, metavar="PATH", help="path to the vocabulary (one token per line)")),
    (("--config_file",), dict(type=str, metavar="PATH", required=True, help="path to the BERT json config")),
    (("--max_pred_length",), dict(type=int, default=512, help="max number of tokens in a sentence")),
]

_FINE_TUNE = [
    (("--hetseq_state_dict",), dict(type=str, default=None, help="checkpoint written by this framework / HetSeq")),
    (("--transformers_state_dict",), dict(type=str, default=None, help="HF transformers state dict")),
    (("--train_file",), dict(type=str, default=None, help="CoNLL-style training file")),
    (("--validation_file",), dict(type=str, default=None, help="CoNLL-style validation file")),
    (("--test_file",), dict(type=str, default=None, help="CoNLL-style test file")),
    (("--extension_file",), dict(type=str, default=None, help="HF datasets loading script (optional)")),
    (("--load_state_dict_strict",), dict(type=eval_bool, default=False, help="strict state-dict loading")),
]

_TASK_ROWS = {
    "bert": _BERT_FILES + [
        (("--num_file",), dict(type=int, default=0, help="number of shard files to use, 0 for all")),
    ],
    "mnist": [],
    "BertForTokenClassification": _BERT_FILES + _FINE_TUNE,
    "BertForELClassification": _BERT_FILES + _FINE_TUNE + [
        (("--root_data_dir",), dict(type=str, default="", help="root of the deep_ed data directory")),
        (("--entities",), dict(type=str, default="RLTD", choices=["RLTD", "4EX", "ALL"], help="entity set")),
        (("--ent_vecs_filename",), dict(type=str, default="", help="pretrained entity embedding tensor (.pt)")),
    ],
}

_OPT_COMMON = [
    (("--max-epoch", "--me"), dict(default=0, type=int, metavar="N", help="force stop training at specified epoch")),
    (("--max-update", "--mu"), dict(default=0, type=int, metavar="N", help="force stop training at specified update")),
    (("--clip-norm",), dict(default=25, type=float, metavar="NORM", help="clip threshold of gradients")),
    (("--update-freq",), dict(default="1", metavar="N1,N2,...,N_K", type=lambda s: eval_str_list(s, type=int),
                              help="update parameters every N_i batches, when in epoch i")),
    (("--lr", "--learning-rate"), dict(default="0.25", type=eval_str_list, metavar="LR_1,LR_2,...,LR_N",
                                       help="learning rate for the first N epochs; all epochs >N using LR_N")),
    (("--min-lr",), dict(default=-1, type=float, metavar="LR", help="stop training when the lr reaches this minimum")),
    (("--use-bmuf",), dict(default=False, action="store_true",
                           help="accepted for script compatibility; BMUF is not implemented (no-op with a warning)")),
]

_OPT_ROWS = {
    "adam": [
        (("--optimizer",), dict(default="adam", type=str, help="optimizer name")),
        (("--adam-betas",), dict(default="(0.9, 0.999)", metavar="B", help="betas for Adam optimizer")),
        (("--adam-eps",), dict(type=float, default=1e-8, metavar="D", help="epsilon for Adam optimizer")),
        (("--weight-decay", "--wd"), dict(default=0.0, type=float, metavar="WD", help="weight decay")),
    ],
    "adadelta": [
        (("--optimizer",), dict(default="adadelta", type=str, help="optimizer name")),
        (("--adadelta_rho",), dict(default=0.9, type=float)),
        (("--adadelta_eps",), dict(default=1e-6, type=float)),
        # the reference spells this flag without the leading "a" (hetseq/options.py:275); keep both
        (("--dadelta_weight_decay", "--adadelta_weight_decay"), dict(default=0.0, type=float, dest="dadelta_weight_decay")),
    ],
}

_SCHED_ROWS = {
    "PolynomialDecayScheduler": [
        (("--lr_scheduler", "--lr-scheduler"), dict(default="PolynomialDecayScheduler", type=str, help="lr scheduler name")),
        (("--force-anneal", "--fa"), dict(type=int, metavar="N", help="force annealing at specified epoch")),
        (("--warmup-updates",), dict(default=0, type=int, metavar="N", help="linear warm-up over the first N updates")),
        (("--end-learning-rate",), dict(default=0.0, type=float)),
        (("--power",), dict(default=1.0, type=float)),
        (("--total-num-update",), dict(default=1000000, type=int)),
    ],
}

_CHECKPOINT = [
    (("--save-dir",), dict(metavar="DIR", default="checkpoints", help="path to save checkpoints")),
    (("--restore-file",), dict(default="checkpoint_last.pt", help="checkpoint to resume from (relative to --save-dir)")),
    (("--reset-dataloader",), dict(action="store_true", help="do not restore the data iterator position")),
    (("--reset-lr-scheduler",), dict(action="store_true", help="do not restore the lr scheduler state")),
    (("--reset-meters",), dict(action="store_true", help="do not restore meters")),
    (("--reset-optimizer",), dict(action="store_true", help="do not restore optimizer state")),
    (("--optimizer-overrides",), dict(default="{}", type=str, metavar="DICT", help="override optimizer args on load")),
    (("--save-interval",), dict(type=int, default=1, metavar="N", help="save a checkpoint every N epochs")),
    (("--save-interval-updates",), dict(type=int, default=0, metavar="N", help="save (and validate) every N updates")),
    (("--keep-interval-updates",), dict(type=int, default=-1, metavar="N", help="keep the last N update checkpoints")),
    (("--keep-last-epochs",), dict(type=int, default=-1, metavar="N", help="keep last N epoch checkpoints")),
    (("--no-save",), dict(action="store_true", help="don't save models or checkpoints")),
    (("--no-epoch-checkpoints",), dict(action="store_true", help="only store last and best checkpoints")),
    (("--no-last-checkpoints",), dict(action="store_true", help="don't store last checkpoints")),
    (("--no-save-optimizer-state",), dict(action="store_true", help="don't save optimizer state")),
    (("--best-checkpoint-metric",), dict(type=str, default="loss", help='metric for "best" checkpoints')),
    (("--maximize-best-checkpoint-metric",), dict(action="store_true", help="larger metric is better")),
]

_ENGINE = [
    (("--precision",), dict(default="tf32", choices=["tf32", "bf16", "fp32"],
                            help="tf32: fp32 storage, TF32 tensor-core math (reference parity); "
                                 "bf16: bf16 activations/weights with fp32 master weights; "
                                 "fp32: fp32 storage and fp32-accurate products (every GEMM as three TF32 products of "
                                 "split operands, ~3.5x the TF32 cost; plain PyTorch fp32 with --no-native / --cpu)")),
    (("--comm",), dict(default="auto", choices=["auto", "p2p", "torch"],
                       help="gradient reduction: hand-written NVLink peer kernels (p2p) or torch.distributed (torch)")),
    (("--p2p-bucket-mb",), dict(type=int, default=100, metavar="MB",
                                help="bucket size of the peer-to-peer reducer (--bucket-cap-mb keeps applying to the "
                                     "torch.distributed reducer); 0: use --bucket-cap-mb")),
    (("--no-native",), dict(action="store_true", help="force the plain-PyTorch model path (CPU/debug only)")),
    (("--cuda-graph",), dict(action=argparse.BooleanOptionalAction, default=None,
                             help="capture the train step (all micro-batches of an update) in a CUDA graph per batch "
                                  "shape and replay it; default: on when training on CUDA")),
    (("--no-sharded-optimizer",), dict(action="store_true",
                                       help="peer-to-peer reducer: all-reduce the gradients and run the optimizer "
                                            "replicated on every rank instead of reduce-scatter + owned-slice update + "
                                            "parameter broadcast")),
    (("--pipelined-update",), dict(action="store_true",
                                   help="sharded optimizer step: park it at the end of the step and issue it (with its "
                                        "parameter broadcast) bucket by bucket under the NEXT forward pass instead of running "
                                        "it exposed; measured +0.4%% at 8 GPUs, -1%% at 2 (profiles/README.md), hence opt-in")),
    (("--checkpoint-activations",), dict(action="store_true", help="recompute encoder activations in backward")),
    (("--pad-to-multiple-of",), dict(type=int, default=0, metavar="N",
                                     help="NER / EL collators: pad the batch width up to a multiple of N (0: longest "
                                          "sequence, as the reference); with --cuda-graph this bounds the number of "
                                          "distinct batch shapes, i.e. of captured graphs")),
    (("--device-batch-assembly",), dict(action=argparse.BooleanOptionalAction, default=None,
                                        help="ship BERT batches as one packed int32 matrix and build the int64 inputs / "
                                             "dense MLM labels with a kernel on the GPU; default: on when training on CUDA")),
    (("--max-predictions-per-seq",), dict(type=int, default=0, metavar="P",
                                          help="upper bound of masked positions per sequence (0: sequence length)")),
    (("--sync-stats",), dict(action="store_true",
                             help="wait for each step's statistics before returning from train_step (default: they reach "
                                  "the meters one step late through a pinned mailbox, so the host never waits for the GPU)")),
    (("--profile",), dict(action="store_true", help="emit NVTX ranges per phase")),
    (("--log-jsonl",), dict(type=str, default=None, metavar="PATH", help="also append per-step stats as JSON lines")),
    (("--fault-inject",), dict(type=str, default=None, metavar="RANK:STEP", help="testing: kill RANK at STEP")),
    (("--max-restarts",), dict(type=int, default=0, metavar="N",
                               help="when a rank of this node dies, stop the others and start all of them again from "
                                    "checkpoint_last.pt, at most N times (use with --save-interval-updates)")),
]


# -------------------------------------------------------------------------------------------------
# public API (names match the reference)
# -------------------------------------------------------------------------------------------------
def add_dataset_args(parser, train: bool = False, gen: bool = False, task: str = "bert"):
    group = parser.add_argument_group("Dataset and data loading")
    _add_rows(group, _DATA_COMMON)
    if train:
        _add_rows(group, _DATA_TRAIN)
        if task not in _TASK_ROWS:
            raise ValueError("unsupported task: {}".format(task))
        parser.add_argument("--task", type=str, default=task)
        parser.add_argument("--data", type=str, help="path including data")
        _add_rows(group, _TASK_ROWS[task])
    return group


def add_distributed_training_args(parser):
    group = parser.add_argument_group("Distributed training")
    _add_rows(group, [
        (("--distributed-world-size",), dict(type=int, metavar="N", default=_default_world_size(),
                                             help="total number of GPUs across all nodes (default: all visible GPUs)")),
        (("--distributed-rank",), dict(default=0, type=int, help="rank of the first GPU of this node")),
        (("--distributed-gpus",), dict(default=4, type=int, help="number of GPUs used on the current node")),
        (("--distributed-backend",), dict(default="nccl", type=str, help="distributed backend (nccl | gloo)")),
        (("--distributed-init-method",), dict(default=None, type=str,
                                              help="tcp://host:port or file:///shared/path rendezvous")),
        (("--device-id", "--local_rank"), dict(default=0, type=int, help="which GPU to use (set automatically)")),
        (("--distributed-no-spawn",), dict(action="store_true", help="do not spawn one process per GPU")),
        (("--ddp-backend",), dict(default="c10d", type=str, choices=["c10d"], help="kept for script compatibility")),
        (("--bucket-cap-mb",), dict(default=25, type=int, metavar="MB", help="gradient bucket size")),
        (("--fix-batches-to-gpus",), dict(action="store_true", help="don't reshuffle batches between GPUs per epoch")),
        (("--find-unused-parameters",), dict(default=False, action="store_true",
                                             help="tolerate parameters that receive no gradient")),
        (("--fast-stat-sync",), dict(default=False, action="store_true",
                                     help="sync a fixed stats vector instead of gathering python objects")),
    ])
    return group


def add_optimization_args(parser, optimizer: str = "adam", lr_scheduler: str = "PolynomialDecayScheduler"):
    group = parser.add_argument_group("Optimization")
    _add_rows(group, _OPT_COMMON)
    if optimizer not in _OPT_ROWS:
        raise ValueError("unsupported optimizer: {}".format(optimizer))
    _add_rows(group, _OPT_ROWS[optimizer])
    if lr_scheduler not in _SCHED_ROWS:
        raise ValueError("unsupported lr_scheduler: {}".format(lr_scheduler))
    _add_rows(group, _SCHED_ROWS[lr_scheduler])
    return group


def add_checkpoint_args(parser):
    group = parser.add_argument_group("Checkpointing")
    _add_rows(group, _CHECKPOINT)
    return group


def add_engine_args(parser):
    group = parser.add_argument_group("B200 engine")
    _add_rows(group, _ENGINE)
    return group


def get_training_parser(task: str = "bert", optimizer: str = "adam", lr_scheduler: str = "PolynomialDecayScheduler"):
    parser = argparse.ArgumentParser(allow_abbrev=False)
    _add_rows(parser, _GENERAL)
    add_dataset_args(parser, train=True, task=task)
    add_distributed_training_args(parser)
    add_optimization_args(parser, optimizer=optimizer, lr_scheduler=lr_scheduler)
    add_checkpoint_args(parser)
    add_engine_args(parser)
    return parser


def preparse(argv: Optional[Sequence[str]] = None):
    """Stage one: find ``--task/--optimizer/--lr-scheduler`` so the full flag set can be built
    (ref: hetseq/train.py:204-212)."""
    pre = argparse.ArgumentParser(add_help=False, allow_abbrev=False)
    # choices come from the flag tables, so an extension that adds its rows (docs/extending.md) is accepted here too
    pre.add_argument("--task", default="bert", choices=list(_TASK_ROWS))
    pre.add_argument("--optimizer", default="adam", choices=list(_OPT_ROWS))
    pre.add_argument("--lr-scheduler", "--lr_scheduler", dest="lr_scheduler", default="PolynomialDecayScheduler",
                     choices=list(_SCHED_ROWS))
    known, _ = pre.parse_known_args(argv)
    return known


def parse_args_and_arch(parser, s=None):
    args = parser.parse_args(s)
    if getattr(args, "max_sentences_valid", None) is None:
        args.max_sentences_valid = getattr(args, "max_sentences", None)
    if getattr(args, "max_tokens_valid", None) is None:
        args.max_tokens_valid = getattr(args, "max_tokens", None)
    # engine switches that default to "on when the step runs on a GPU"
    for name in ("cuda_graph", "device_batch_assembly"):
        if hasattr(args, name) and getattr(args, name) is None:
            setattr(args, name, _on_cuda(args) and (name == "cuda_graph" or getattr(args, "task", "bert") == "bert"))
    return args


def _on_cuda(args) -> bool:
    if getattr(args, "cpu", False) or getattr(args, "no_native", False):
        return False
    try:
        import torch
        return bool(torch.cuda.is_available())
    except Exception:
        return False


def parse_training_args(argv: Optional[Sequence[str]] = None):
    """Both stages in one call: returns the fully parsed ``argparse.Namespace``."""
    sel = preparse(argv)
    parser = get_training_parser(task=sel.task, optimizer=sel.optimizer, lr_scheduler=sel.lr_scheduler)
    return parse_args_and_arch(parser, argv)
