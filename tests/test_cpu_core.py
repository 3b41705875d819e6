"""CPU unit tests: flags, meters / log format, batcher, iterators, LR schedule, optimizers, checkpoints.
Where the reference implementation can be imported (read-only mount at /root/reference, missing third-party
modules stubbed), it is used as the oracle; otherwise the test falls back to closed-form expectations."""
import glob
import math
import os
import re
import shlex
import sys
import types

import numpy as np
import pytest
import torch

REF = "/root/reference"


def _ref_module(name):
    """Import hetseq.<name> from the read-only reference with h5py/boto3 stubbed; skip if unavailable."""
    if not os.path.isdir(os.path.join(REF, "hetseq")):
        pytest.skip("reference mount not present")
    for stub in ("h5py", "boto3", "botocore", "botocore.exceptions"):
        if stub not in sys.modules:
            m = types.ModuleType(stub)
            if stub == "botocore.exceptions":
                m.ClientError = type("ClientError", (Exception,), {})
            sys.modules[stub] = m
    if REF not in sys.path:
        sys.path.append(REF)
        sys.path.append(os.path.join(REF, "hetseq"))
    import importlib
    try:
        return importlib.import_module("hetseq." + name)
    except Exception as e:  # pragma: no cover
        pytest.skip("reference module hetseq.%s not importable here: %r" % (name, e))


# ---------------------------------------------------------------------------------------------------
# options
# ---------------------------------------------------------------------------------------------------
def _script_commands():
    cmds = []
    for path in glob.glob(os.path.join(REF, "STORE_RUN_FILE", "**", "*.sh"), recursive=True):
        text = open(path, errors="ignore").read().replace("\\\n", " ")
        for line in text.splitlines():
            if "train.py" in line and "--task" in line:
                cmds.append((path, line))
    return cmds


def test_every_shipped_training_command_parses():
    from hetseq_9cme_b200 import options
    cmds = _script_commands()
    if not cmds:
        pytest.skip("no reference scripts available")
    parsed = 0
    for path, line in cmds:
        line = re.sub(r"\$\{?\w+\}?", "X", line)         # shell variables -> placeholder
        tail = re.split(r"\s(?:\||>|2>&1|&)\s?", line.split("train.py", 1)[1])[0]     # drop pipes / redirects
        toks = shlex.split(tail, comments=True)
        args = options.parse_training_args(toks)
        assert args.task in options.TASKS
        parsed += 1
    assert parsed >= 10


def test_own_launch_scripts_parse_and_cover_the_reference_layouts():
    """scripts/ (tools/gen_run_scripts.py) must offer every job layout of the reference's STORE_RUN_FILE tree, and every
    command in it must be accepted by our parser with consistent ranks."""
    from hetseq_9cme_b200 import options
    root = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "scripts")
    for layout in ("node1gpu4", "node2gpu8", "node2gpu8_het", "node4gpu16", "node4gpu16_het", "node8gpu32_het"):
        assert os.path.exists(os.path.join(root, "Train_mnist", layout, "main.sh")), layout
        assert os.path.exists(os.path.join(root, "Train_mnist", layout, "eval.sh")), layout
    n = 0
    for path in glob.glob(os.path.join(root, "**", "*.sh"), recursive=True):
        text = open(path).read().replace("\\\n", " ")
        for line in text.splitlines():
            if "train.py" not in line or "--task" not in line:
                continue
            line = re.sub(r"\$\{[^}]*\{[^}]*\}[^}]*\}|\$\{[^}]*\}", "1", line)      # ${A:-$B/c} and ${A}
            tail = re.split(r"\s(?:\||>|2>&1|&)\s?", line.split("train.py", 1)[1])[0]
            args = options.parse_training_args(shlex.split(tail, comments=True))
            assert args.task in options.TASKS
            if "--distributed-rank" in tail:
                assert args.distributed_rank + args.distributed_gpus <= args.distributed_world_size, path
            n += 1
    assert n >= 25


def test_flag_defaults_match_reference_table():
    from hetseq_9cme_b200 import options
    a = options.parse_training_args(["--task", "bert", "--config_file", "c.json", "--max-sentences", "8"])
    assert (a.seed, a.clip_norm, a.bucket_cap_mb, a.distributed_gpus, a.distributed_backend) == (19940802, 25, 25, 4, "nccl")
    assert a.lr == [0.25] and a.update_freq == [1] and a.adam_betas == "(0.9, 0.999)" and a.adam_eps == 1e-8
    assert a.restore_file == "checkpoint_last.pt" and a.save_dir == "checkpoints" and a.max_pred_length == 512
    assert a.max_sentences_valid == 8 and a.total_num_update == 1000000 and a.power == 1.0
    b = options.parse_training_args(["--task", "mnist", "--optimizer", "adadelta", "--max-sentences", "4", "--lr", "1.0,0.5",
                                     "--update-freq", "2,1"])
    assert b.lr == [1.0, 0.5] and b.update_freq == [2, 1] and b.adadelta_rho == 0.9 and b.adadelta_eps == 1e-6
    assert options.eval_str_list("[1,2]", int) == [1, 2] and options.eval_str_list(3) == [3.0]
    assert options.eval_bool("True") is True and options.eval_bool(None, True) is True


# ---------------------------------------------------------------------------------------------------
# meters / progress bar
# ---------------------------------------------------------------------------------------------------
def test_log_line_format(capsys):
    from hetseq_9cme_b200 import meters, progress_bar
    loss = meters.AverageMeter(); loss.update(0.044, 32)
    ups = meters.TimeMeter(); ups.n = 3; ups.start -= 1.0
    sw = meters.StopwatchMeter(); sw.sum = 1.0718
    stats = {"loss": loss, "ppl": 1.03, "ups": ups, "num_updates": 2, "lr": 9.9996e-05, "train_wall": sw}
    bar = progress_bar.simple_progress_bar(list(range(3)), epoch=1, log_interval=1)
    for _ in bar:
        bar.log(stats)
    out = capsys.readouterr().out.splitlines()
    assert out[0].startswith("| epoch 001:      1 / 3 loss=0.044, ppl=1.03, ups=")
    assert "num_updates=2, lr=9.9996e-05" in out[0] and out[0].endswith("train_wall=1.0718")
    assert progress_bar.format_stat(sw) == "1.0718" and progress_bar.format_stat(loss) == "0.044"


def test_meters_roundtrip():
    from hetseq_9cme_b200 import meters
    m = meters.AverageMeter(); m.update(2.0, 3); m.update(4.0, 1)
    assert m.avg == pytest.approx(2.5) and m.count == 4
    n = meters.AverageMeter(); n.load_state_dict(m.state_dict())
    assert n.avg == m.avg
    s = meters.StopwatchMeter(); s.start(); s.stop(); assert s.n == 1 and s.sum >= 0


# ---------------------------------------------------------------------------------------------------
# batching and iterators
# ---------------------------------------------------------------------------------------------------
def test_batch_by_size_matches_reference_cython():
    from hetseq_9cme_b200.data import data_utils
    ref_fast = None
    base = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline")
    if os.path.isdir(os.path.join(base, "_ref", "hetseq")):
        # the package import pulls in h5py / boto3, which the image lacks: the baseline's stand-ins provide them
        sys.path[:0] = [os.path.join(base, "_ref"), os.path.join(base, "shims")]
        try:
            from hetseq.data.data_utils_fast import batch_by_size_fast as ref_fast  # built by the baseline install
        finally:
            del sys.path[:2]
        assert ref_fast is not None
    rng = np.random.RandomState(0)
    for _ in range(60):
        n = int(rng.randint(1, 300))
        sizes = rng.randint(1, 60, size=n).astype(np.int64)
        idx = rng.permutation(n).astype(np.int64)
        mt = int(rng.choice([60, 128, 4096])); ms = int(rng.choice([1, 4, 32, 10 ** 9])); mult = int(rng.choice([1, 2, 8]))
        got = data_utils.batch_by_size(idx, sizes=sizes, max_tokens=mt, max_sentences=ms, required_batch_size_multiple=mult)
        assert sorted(i for b in got for i in b) == sorted(idx.tolist())
        if ref_fast is not None:
            want = ref_fast(idx, lambda i: int(sizes[i]), mt, ms, mult)
            assert got == [list(map(int, b)) for b in want]
        assert all(len(b) <= ms for b in got)
    # BERT case: constant size, contiguous chunks
    assert data_utils.batch_by_size(np.arange(70), sizes=np.full(70, 512), max_sentences=32) == [list(range(32)), list(range(32, 64)), list(range(64, 70))]
    with pytest.raises(AssertionError):
        data_utils.batch_by_size(np.arange(3), sizes=np.array([5, 500, 5]), max_tokens=100)


class _ToyDataset(torch.utils.data.Dataset):
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return torch.tensor([i])

    def collater(self, samples):
        return None if len(samples) == 0 else torch.stack(samples)

    def set_epoch(self, e):
        pass


def test_epoch_iterator_sharding_shuffle_resume():
    from hetseq_9cme_b200.data import iterators
    ds = _ToyDataset(45)
    batches = [list(range(i, min(45, i + 4))) for i in range(0, 45, 4)]        # 12 batches
    world = 5
    its = [iterators.EpochBatchIterator(ds, ds.collater, batches, seed=7, num_shards=world, shard_id=r) for r in range(world)]
    per_rank = [it.batches_for_epoch(1) for it in its]
    assert len({len(p) for p in per_rank}) == 1 and len(per_rank[0]) == math.ceil(12 / world)    # lock-step
    flat = [tuple(b) for p in per_rank for b in p if b]
    assert sorted(flat) == sorted(tuple(b) for b in batches)                    # disjoint cover
    assert sum(1 for p in per_rank for b in p if not b) == world * 3 - 12        # [] padding
    assert its[0].batches_for_epoch(1) == its[0].batches_for_epoch(1)
    assert its[0].batches_for_epoch(1) != its[0].batches_for_epoch(2)           # reshuffled per epoch
    # order equals numpy legacy shuffle with seed + epoch (what the reference does)
    state = np.random.get_state(); np.random.seed(7 + 1); ref = list(batches); np.random.shuffle(ref); np.random.set_state(state)
    assert per_rank[0] == ref[0::world]
    # iterate, stop mid-epoch, resume
    it = iterators.EpochBatchIterator(ds, ds.collater, batches, seed=7, num_shards=1, shard_id=0, num_workers=2)
    e = it.next_epoch_itr(shuffle=True)
    first = [next(e).flatten().tolist() for _ in range(5)]
    st = it.state_dict()
    assert st == {"epoch": 1, "iterations_in_epoch": 5, "shuffle": True}
    it2 = iterators.EpochBatchIterator(ds, ds.collater, batches, seed=7)
    it2.load_state_dict(st)
    rest = [b.flatten().tolist() for b in it2.next_epoch_itr(shuffle=True)]
    full = [b.flatten().tolist() for b in iterators.EpochBatchIterator(ds, ds.collater, batches, seed=7).next_epoch_itr()]
    assert first + rest == full and it2.epoch == 1
    g = iterators.GroupedIterator(iterators.CountingIterator(list(range(7))), 3)
    assert len(g) == 3 and [c for c in g] == [[0, 1, 2], [3, 4, 5], [6]]
    sh = iterators.ShardedIterator(list(range(7)), 3, 2, fill_value=-1)
    assert list(sh) == [2, 5, -1]


# ---------------------------------------------------------------------------------------------------
# lr schedule / optimizers
# ---------------------------------------------------------------------------------------------------
def _opt_args(**kw):
    from argparse import Namespace
    base = dict(lr=[1e-3], adam_betas="(0.9, 0.999)", adam_eps=1e-8, weight_decay=0.01, warmup_updates=4, end_learning_rate=0.0,
                power=1.0, total_num_update=20, force_anneal=None, adadelta_rho=0.9, adadelta_eps=1e-6, dadelta_weight_decay=0.0,
                optimizer="adam", lr_scheduler="PolynomialDecayScheduler")
    base.update(kw)
    return Namespace(**base)


def test_polynomial_schedule_values_and_reference_parity():
    from hetseq_9cme_b200 import lr_scheduler, optim
    args = _opt_args()
    p = torch.nn.Parameter(torch.zeros(3))
    o = optim._Adam(args, [p])
    s = lr_scheduler.PolynomialDecayScheduler(args, o)
    got = [s.step_update(n) for n in range(0, 24)]
    assert got[0] == 0.0                                  # lr = 0 on the very first update (reference quirk)
    assert got[2] == pytest.approx(0.5e-3) and got[4] == pytest.approx(1e-3)
    assert got[12] == pytest.approx(1e-3 * (1 - 8 / 16)) and got[20] == 0.0 and got[23] == 0.0
    ref_sched = _ref_module("lr_scheduler")
    ref_optim = _ref_module("optim")
    q = torch.nn.Parameter(torch.zeros(3))
    ro = ref_optim._Adam(_opt_args(), [q])
    rs = ref_sched.PolynomialDecayScheduler(_opt_args(), ro)
    assert got == [rs.step_update(n) for n in range(0, 24)]


@pytest.mark.parametrize("name", ["adam", "adadelta"])
def test_optimizers_match_reference_bit_for_bit(name):
    from hetseq_9cme_b200 import optim
    ref_optim = _ref_module("optim")
    torch.manual_seed(0)
    shapes = [(7, 5), (11,), (3, 2, 2)]
    mine = [torch.nn.Parameter(torch.randn(s)) for s in shapes]
    theirs = [torch.nn.Parameter(p.detach().clone()) for p in mine]
    args = _opt_args(optimizer=name, lr=[0.05], weight_decay=0.01, dadelta_weight_decay=0.01)
    a = (optim._Adam if name == "adam" else optim._Adadelta)(args, mine)
    b = (ref_optim._Adam if name == "adam" else ref_optim._Adadelta)(args, theirs)
    for step in range(5):
        for p, q in zip(mine, theirs):
            g = torch.randn_like(p)
            p.grad = g.clone(); q.grad = g.clone()
        a.multiply_grads(0.5); b.multiply_grads(0.5)
        na, nb = a.clip_grad_norm(0.7), b.clip_grad_norm(0.7)
        assert float(na) == pytest.approx(float(nb), rel=1e-6)
        a.step(); b.step()
        for p, q in zip(mine, theirs):
            assert torch.allclose(p, q, rtol=1e-6, atol=1e-7), (name, step)
    assert set(a.state_dict()["state"][0].keys()) == set(b.state_dict()["state"][0].keys())
    assert type(a).__name__ == type(b).__name__


def test_arena_optimizer_equals_per_tensor_path_on_cpu():
    from hetseq_9cme_b200 import optim
    from hetseq_9cme_b200.parallel.arena import ParamArena
    torch.manual_seed(1)
    model_a = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Linear(5, 3))
    model_b = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Linear(5, 3))
    model_b.load_state_dict(model_a.state_dict())
    arena = ParamArena(list(model_a.parameters()), device="cpu")
    args = _opt_args(lr=[0.01])
    oa = optim._Adam(args, arena.params); oa.attach_arena(arena)
    ob = optim._Adam(args, list(model_b.parameters()))
    x = torch.randn(4, 6)
    for _ in range(3):
        model_a(x).pow(2).sum().backward(); model_b(x).pow(2).sum().backward()
        assert float(oa.clip_grad_norm(0.5)) == pytest.approx(float(ob.clip_grad_norm(0.5)), rel=1e-5)
        oa.step(); ob.step(); oa.zero_grad(); ob.zero_grad()
    for p, q in zip(model_a.parameters(), model_b.parameters()):
        assert torch.allclose(p, q, atol=1e-6)
    assert all(p.grad is not None and p.grad.abs().sum() == 0 for p in model_a.parameters())   # views survive zero_grad
    sd = oa.state_dict()
    assert sd["state"][0]["exp_avg"].shape == (5, 6)
    oa.load_state_dict(sd)
    assert oa.optimizer.state[arena.params[0]]["exp_avg"].data_ptr() == oa.optimizer._flat_state["exp_avg"].data_ptr()


# ---------------------------------------------------------------------------------------------------
# checkpoints
# ---------------------------------------------------------------------------------------------------
def test_checkpoint_schema_and_resume(tmp_path):
    from hetseq_9cme_b200 import train, checkpoint_utils
    from hetseq_9cme_b200.data import synthetic
    d = str(tmp_path)
    synthetic.make_mnist(d, 256, 64)
    common = ["--task", "mnist", "--optimizer", "adadelta", "--data", d, "--cpu", "--max-sentences", "32", "--lr", "1.0",
              "--valid-subset", "test", "--save-dir", d + "/ckpt", "--distributed-world-size", "1", "--num-workers", "0",
              "--clip-norm", "100", "--log-format", "none", "--save-interval-updates", "5", "--keep-interval-updates", "1"]
    c1 = train.main(__import__("hetseq_9cme_b200.options", fromlist=["x"]).parse_training_args(common + ["--max-epoch", "1"]))
    files = sorted(os.listdir(d + "/ckpt"))
    assert "checkpoint1.pt" in files and "checkpoint_last.pt" in files and "checkpoint_best.pt" in files
    assert sum(1 for f in files if re.fullmatch(r"checkpoint_\d+_\d+\.pt", f)) == 1            # pruned to the newest
    state = checkpoint_utils.load_checkpoint_to_cpu(d + "/ckpt/checkpoint_last.pt")
    assert set(state.keys()) == {"args", "model", "optimizer_history", "extra_state", "last_optimizer_state"}
    assert state["optimizer_history"][-1]["optimizer_name"] == "_Adadelta"
    assert state["optimizer_history"][-1]["num_updates"] == 8
    assert state["extra_state"]["train_iterator"]["epoch"] == 1 and "train_meters" in state["extra_state"]
    assert list(state["model"].keys())[0] == "conv1.weight" and state["model"]["fc2.bias"].dtype == torch.float32
    assert set(state["last_optimizer_state"]["state"][0].keys()) == {"step", "square_avg", "acc_delta"}
    # resume: continues at update 8, epoch 2
    c2 = train.main(__import__("hetseq_9cme_b200.options", fromlist=["x"]).parse_training_args(common + ["--max-epoch", "2"]))
    assert c2.get_num_updates() == 16
    # a run that never stopped reaches the same weights
    import shutil
    shutil.rmtree(d + "/ckpt")
    c3 = train.main(__import__("hetseq_9cme_b200.options", fromlist=["x"]).parse_training_args(common + ["--max-epoch", "2"]))
    for (k, a), (_, b) in zip(c2.get_model().state_dict().items(), c3.get_model().state_dict().items()):
        assert torch.allclose(a, b, atol=1e-6), k


def test_reference_style_checkpoint_with_empty_extra_state_loads(tmp_path):
    """The reference writes ``extra_state = {}`` (hetseq/checkpoint_utils.py:204); such files must still load."""
    from argparse import Namespace
    from hetseq_9cme_b200 import options, tasks
    from hetseq_9cme_b200.controller import Controller
    from hetseq_9cme_b200.data import synthetic
    d = str(tmp_path)
    synthetic.make_mnist(d, 64, 64)
    args = options.parse_training_args(["--task", "mnist", "--optimizer", "adadelta", "--data", d, "--cpu", "--max-sentences", "32",
                                        "--lr", "1.0", "--save-dir", d, "--distributed-world-size", "1"])
    net = tasks.MNISTNet()
    torch.save({"args": Namespace(x=1), "model": net.state_dict(),
                "optimizer_history": [{"optimizer_name": "_Adadelta", "lr_scheduler_state": {"best": None}, "num_updates": 3}],
                "extra_state": {}}, d + "/ref.pt")
    task = tasks.MNISTTask.setup_task(args)
    c = Controller(args, task, tasks.MNISTNet())
    extra = c.load_checkpoint(d + "/ref.pt")
    assert extra == {}
    for (k, a), (_, b) in zip(c.get_model().state_dict().items(), net.state_dict().items()):
        assert torch.equal(a, b), k


def test_public_api_surface_of_the_reference_is_present():
    """Every module / class / function name a HetSeq user imports (taken from the reference's
    `hetseq/` tree) exists under the same name here."""
    import importlib
    surface = {
        "train": "cli_main distributed_main main train get_training_stats",
        "options": "get_training_parser add_dataset_args add_distributed_training_args add_optimization_args "
                   "add_checkpoint_args eval_str_list eval_bool parse_args_and_arch",
        "controller": "Controller",
        "distributed_utils": "distributed_init is_master suppress_output get_rank get_world_size get_default_group "
                             "all_reduce all_gather_list",
        "checkpoint_utils": "save_checkpoint load_checkpoint load_checkpoint_to_cpu checkpoint_paths torch_persistent_save "
                            "convert_state_dict_type save_state verify_checkpoint_directory",
        "optim": "_Optimizer _Adam _Adadelta Adam Adadelta",
        "lr_scheduler": "_LRScheduler PolynomialDecayScheduler",
        "meters": "AverageMeter TimeMeter StopwatchMeter",
        "progress_bar": "build_progress_bar progress_bar noop_progress_bar simple_progress_bar format_stat",
        "tasks": "Task LanguageModelingTask MNISTTask MNISTNet BertForTokenClassificationTask BertForELClassificationTask",
        "data": "BertH5pyData ConBertH5pyData MNISTDataset BertNerDataset BertELDataset",
        "data.data_utils": "numpy_seed batch_by_size",
        "data.iterators": "CountingIterator EpochBatchIterator GroupedIterator ShardedIterator",
        "data_collator": "YD_DataCollatorForTokenClassification YD_DataCollatorForELClassification",
        "bert_modeling": "BertConfig BertLayerNorm LinearActivation BertModel BertForPreTraining BertForMaskedLM "
                         "BertForNextSentencePrediction BertForSequenceClassification BertForMultipleChoice "
                         "BertForTokenClassification BertForQuestionAnswering",
        "model": "BertForELClassification",
        "eval_mnist": "main",
        "eval_bert_fine_tuning_ner": "main",
    }
    missing = []
    for mod, names in surface.items():
        m = importlib.import_module("hetseq_9cme_b200." + mod)
        missing += ["%s.%s" % (mod, n) for n in names.split() if not hasattr(m, n)]
    assert not missing, missing
    assert hasattr(importlib.import_module("hetseq_9cme_b200.controller").Controller, "valid_step")


def test_sharded_step_protocol_has_no_race_in_any_interleaving():
    """Model check of the cross-rank protocol of the sharded data-parallel step (tools/protocol_model.py: stand-in for
    a racecheck across processes): every interleaving of 3 ranks x 2 buckets x 2 steps and of 2 ranks x 3
    buckets x 3 steps is free of stale reads, premature overwrites and deadlocks — and dropping ANY one synchronisation
    element (the wait for the side stream, the announce-wait of the reduce-scatter or the tail, the tail's 'done
    reading' round, the completion round of the sharded Adam, the per-bucket wait of the pipelined update) is caught."""
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tools"))
    try:
        import protocol_model as pm
    finally:
        sys.path.pop(0)
    for pipelined in (False, True):                      # update at the end of the step / --pipelined-update
        for shape in ((3, 2, 2), (2, 3, 3)):
            violation, states = pm.check(*shape, pipelined=pipelined)
            assert violation is None and states > 500, (shape, pipelined, violation, states)
        for drop in (pm.DROPS_PIPELINED if pipelined else pm.DROPS):
            violation, _ = pm.check(2, 2, 3, drop=drop, pipelined=pipelined)
            assert violation is not None, (drop, pipelined)
    # inside the one-shot kernel: fine while the whole grid is co-resident, a certain deadlock beyond (hence the
    # kOneShotMaxBlocks guard in b200_p2p_allreduce)
    assert pm.check_one_shot(2, 3, 3)[0] is None and pm.check_one_shot(3, 2, 2)[0] is None
    assert "deadlock" in pm.check_one_shot(2, 4, 3)[0]


def test_jsonl_sink_mirrors_the_training_log(tmp_path):
    """--log-jsonl: every logged step and the end-of-epoch summaries as JSON lines with the reference's stat names
    (SURVEY §5.5 "optional JSONL sink"), next to the unchanged text log."""
    import json
    from hetseq_9cme_b200 import options, train
    from hetseq_9cme_b200.data import synthetic
    d = str(tmp_path)
    synthetic.make_mnist(d, 16 * 4, 32)
    path = os.path.join(d, "log.jsonl")
    argv = ["--task", "mnist", "--optimizer", "adadelta", "--data", d, "--cpu", "--max-sentences", "16", "--lr", "1.0",
            "--valid-subset", "test", "--save-dir", os.path.join(d, "ckpt"), "--distributed-world-size", "1", "--num-workers", "0",
            "--clip-norm", "100", "--log-format", "simple", "--log-interval", "1", "--max-epoch", "1", "--no-save", "--seed", "3",
            "--log-jsonl", path]
    train.main(options.parse_training_args(argv))
    recs = [json.loads(line) for line in open(path)]
    steps = [r for r in recs if "update" in r]
    assert len(steps) >= 3 and all({"loss", "ppl", "wps", "ups", "num_updates", "lr", "gnorm", "clip"} <= set(r) for r in steps)
    assert [r["num_updates"] for r in steps] == sorted(r["num_updates"] for r in steps)
    assert any(r.get("tag") == "train" for r in recs) and any("valid" in str(r.get("tag")) for r in recs)


def _mnist_cli(tmp, name, extra, data):
    """One `train.main` run on synthetic MNIST; returns (controller, captured stdout, save dir)."""
    import contextlib
    import io
    from hetseq_9cme_b200 import options, train
    save = os.path.join(tmp, name)
    argv = ["--task", "mnist", "--optimizer", "adadelta", "--data", data, "--cpu", "--max-sentences", "16", "--lr", "1.0",
            "--valid-subset", "test", "--save-dir", save, "--distributed-world-size", "1", "--num-workers", "0",
            "--clip-norm", "100", "--log-format", "simple", "--log-interval", "100", "--seed", "3"] + extra
    buf = io.StringIO()
    with contextlib.redirect_stdout(buf):
        c = train.main(options.parse_training_args(argv))
    return c, buf.getvalue(), save


def test_checkpoint_and_schedule_flags_behave_like_the_reference(tmp_path):
    """The resume / checkpoint-policy / stopping flags of hetseq/options.py:238-349, end to end through train.main:
    --reset-* and --optimizer-overrides on resume, --max-update, --min-lr, --validate-interval, --disable-validation,
    --no-epoch-checkpoints, --no-last-checkpoints, --no-save-optimizer-state, --keep-last-epochs, a per-epoch
    --update-freq list, --max-tokens batching."""
    from hetseq_9cme_b200.data import synthetic
    tmp = str(tmp_path)
    data = os.path.join(tmp, "data")
    synthetic.make_mnist(data, 16 * 6, 32)                      # 6 batches of 16 per epoch

    c, out, save = _mnist_cli(tmp, "a", ["--max-epoch", "2"], data)
    assert c.get_num_updates() == 12
    assert sorted(os.listdir(save)) == ["checkpoint1.pt", "checkpoint2.pt", "checkpoint_best.pt", "checkpoint_last.pt"]
    # plain resume continues at epoch 3; with every --reset flag the weights are kept but counting starts over
    c, out, _ = _mnist_cli(tmp, "a", ["--max-epoch", "3"], data)
    assert "(epoch 2 @ 12 updates)" in out and c.get_num_updates() == 18
    c, out, _ = _mnist_cli(tmp, "a", ["--max-epoch", "1", "--reset-optimizer", "--reset-lr-scheduler", "--reset-meters",
                                      "--reset-dataloader"], data)
    assert "@ 0 updates)" in out and c.get_num_updates() == 6
    c, out, _ = _mnist_cli(tmp, "a", ["--max-epoch", "2", "--optimizer-overrides", "{'rho': 0.5}"], data)
    assert c.optimizer.optimizer.param_groups[0]["rho"] == 0.5

    c, out, save = _mnist_cli(tmp, "b", ["--max-epoch", "10", "--max-update", "5"], data)
    assert c.get_num_updates() == 5
    c, out, _ = _mnist_cli(tmp, "c", ["--max-epoch", "5", "--min-lr", "2.0"], data)
    assert c.get_num_updates() == 0                                    # lr 1.0 is already below --min-lr
    c, out, _ = _mnist_cli(tmp, "d", ["--max-epoch", "2", "--validate-interval", "2"], data)
    assert out.count("valid on 'test'") == 1 and "epoch 002 | valid" in out
    c, out, _ = _mnist_cli(tmp, "e", ["--max-epoch", "1", "--disable-validation"], data)
    assert "valid on" not in out

    c, out, save = _mnist_cli(tmp, "f", ["--max-epoch", "2", "--no-epoch-checkpoints", "--no-save-optimizer-state"], data)
    assert sorted(os.listdir(save)) == ["checkpoint_best.pt", "checkpoint_last.pt"]
    assert "last_optimizer_state" not in torch.load(os.path.join(save, "checkpoint_last.pt"), weights_only=False)
    c, out, save = _mnist_cli(tmp, "g", ["--max-epoch", "2", "--no-last-checkpoints"], data)
    assert "checkpoint_last.pt" not in os.listdir(save) and "checkpoint2.pt" in os.listdir(save)
    c, out, save = _mnist_cli(tmp, "h", ["--max-epoch", "3", "--keep-last-epochs", "1"], data)
    assert sorted(os.listdir(save)) == ["checkpoint3.pt", "checkpoint_best.pt", "checkpoint_last.pt"]

    c, out, _ = _mnist_cli(tmp, "i", ["--max-epoch", "2", "--update-freq", "2,1"], data)
    assert c.get_num_updates() == 3 + 6                                # epoch 1 accumulates two batches per update
    c, out, _ = _mnist_cli(tmp, "j", ["--max-epoch", "1", "--max-tokens", "8"], data)
    assert c.get_num_updates() == 12                                   # one "token" per image: batches of 8


def test_extension_contract_new_task_dataset_model_scheduler(tmp_path):
    """docs/extending.md (the reference's docs/source/extending.rst): a task, a dataset, a model and an LR scheduler
    written OUTSIDE the package, registered through the public tables, trained by the unmodified `train.main` — arena,
    fused step, checkpoints and all."""
    import torch.nn as nn
    from hetseq_9cme_b200 import lr_scheduler, options, tasks, train
    from hetseq_9cme_b200.tasks import Task

    class LineDataset(torch.utils.data.Dataset):
        """y = 3 x0 - 2 x1 + 1 with a little noise."""

        def __init__(self, n, seed):
            g = torch.Generator().manual_seed(seed)
            self.x = torch.randn(n, 2, generator=g)
            self.y = 3 * self.x[:, 0] - 2 * self.x[:, 1] + 1 + 0.01 * torch.randn(n, generator=g)

        def __len__(self):
            return len(self.x)

        def __getitem__(self, i):
            return self.x[i], self.y[i]

        def ordered_indices(self):
            return np.arange(len(self))

        def num_tokens(self, i):
            return 1

        def collater(self, samples):
            if len(samples) == 0:
                return None
            return [torch.stack([s[0] for s in samples]), torch.stack([s[1] for s in samples])]

        def set_epoch(self, epoch):
            pass

    class LineModel(nn.Module):
        def __init__(self):
            super().__init__()
            self.lin = nn.Linear(2, 1)

        def forward(self, x, y):
            return ((self.lin(x).squeeze(-1) - y) ** 2).mean()

    class LineTask(Task):
        def load_dataset(self, split, **kwargs):
            self.datasets[split] = LineDataset(self.args.line_samples, seed=0 if split == "train" else 1)

        def build_model(self, args):
            return LineModel()

        def _sample_size(self, sample):
            return 1                                   # the loss is already a mean

    class HalvingScheduler(lr_scheduler._LRScheduler):
        """lr = base / 2**(epoch // 4)."""

        def step(self, epoch, val_loss=None):
            super().step(epoch, val_loss)
            self.optimizer.set_lr(self.args.lr[0] / (2 ** (epoch // 4)) * self.args.halving_scale)
            return self.optimizer.get_lr()

    tasks.TASK_REGISTRY["line"] = LineTask
    options._TASK_ROWS["line"] = [(("--line-samples",), dict(type=int, default=64))]
    lr_scheduler.SCHEDULER_REGISTRY["halving"] = HalvingScheduler
    options._SCHED_ROWS["halving"] = [(("--lr_scheduler", "--lr-scheduler"), dict(default="halving", type=str)),
                                      (("--halving-scale",), dict(type=float, default=1.0))]
    try:
        argv = ["--task", "line", "--line-samples", "128", "--optimizer", "adam", "--lr-scheduler", "halving", "--halving-scale", "1.0",
                "--lr", "0.2", "--cpu", "--max-sentences", "16", "--valid-subset", "valid", "--save-dir", str(tmp_path / "ck"),
                "--distributed-world-size", "1", "--num-workers", "0", "--log-format", "none", "--max-epoch", "8", "--clip-norm", "0",
                "--no-epoch-checkpoints"]
        args = options.parse_training_args(argv)
        assert args.line_samples == 128 and args.halving_scale == 1.0
        c = train.main(args)
        w = c.get_model().lin.weight.detach().flatten().tolist()
        assert abs(w[0] - 3) < 0.2 and abs(w[1] + 2) < 0.2 and abs(float(c.get_model().lin.bias) - 1) < 0.2
        assert c.get_num_updates() == 8 * 8 and c.get_lr() == pytest.approx(0.2 / 2 ** 2)
        assert os.path.exists(str(tmp_path / "ck" / "checkpoint_last.pt")) and c.get_meter("valid_loss").avg < 0.1
    finally:
        tasks.TASK_REGISTRY.pop("line", None)
        options._TASK_ROWS.pop("line", None)
        lr_scheduler.SCHEDULER_REGISTRY.pop("halving", None)
        options._SCHED_ROWS.pop("halving", None)


def test_our_mnist_checkpoint_loads_into_the_reference_evaluator(tmp_path):
    """'Checkpoints go both ways': a checkpoint written by this framework loads strictly into the model class of the
    reference's own evaluation script (hetseq/eval_mnist.py:9-36, :48-51) and gives the same outputs."""
    pytest.importorskip("torchvision")
    ref_path = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "baseline", "_ref", "hetseq", "eval_mnist.py")
    if not os.path.exists(ref_path):
        ref_path = "/root/reference/hetseq/eval_mnist.py"
    if not os.path.exists(ref_path):
        pytest.skip("reference not available")
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_eval_mnist", ref_path)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    from hetseq_9cme_b200.data import synthetic
    data = os.path.join(str(tmp_path), "data")
    synthetic.make_mnist(data, 16 * 4, 32)
    c, _, save = _mnist_cli(str(tmp_path), "ck", ["--max-epoch", "1"], data)
    ckpt = torch.load(os.path.join(save, "checkpoint_last.pt"), map_location="cpu", weights_only=False)
    net = ref.MNISTNet()
    net.load_state_dict(ckpt["model"], strict=True)
    net.eval()
    ours = c.get_model().eval()
    x, y = torch.randn(5, 1, 28, 28), torch.randint(0, 10, (5,))
    with torch.no_grad():
        out_ref, loss_ref = net(x, y, eval=True)
        loss_ours = ours(x, y)
    assert torch.allclose(loss_ref, loss_ours, atol=1e-6) and out_ref.shape == (5, 10)
