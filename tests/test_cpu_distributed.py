"""Two-process CPU/gloo runs (BASELINE config #1 — something the reference itself cannot do: its stat sync and DDP
wrapper are CUDA-only, hetseq/controller.py:81,295)."""
import os
import sys

import pytest
import torch
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, data_dir, out_dir, fast):
    sys.path.insert(0, ROOT)
    from hetseq_9cme_b200 import options, train
    argv = ["--task", "mnist", "--optimizer", "adadelta", "--data", data_dir, "--cpu", "--max-sentences", "16", "--lr", "1.0",
            "--valid-subset", "test", "--save-dir", os.path.join(out_dir, "ckpt"), "--distributed-world-size", str(world),
            "--distributed-backend", "gloo", "--distributed-init-method", "tcp://127.0.0.1:%d" % port,
            "--distributed-rank", str(rank), "--num-workers", "0", "--clip-norm", "100", "--log-format", "none",
            "--max-epoch", "1", "--no-save", "--seed", "3"] + (["--fast-stat-sync"] if fast else [])
    args = options.parse_training_args(argv)
    args.device_id = 0
    c = train.main(args, init_distributed=True)
    torch.save({k: v.clone() for k, v in c.get_model().state_dict().items()}, os.path.join(out_dir, "rank%d.pt" % rank))
    torch.save({"updates": c.get_num_updates(), "loss": c.get_meter("train_loss").avg}, os.path.join(out_dir, "meta%d.pt" % rank))


@pytest.mark.parametrize("fast", [False, True])
def test_mnist_gloo_world2_ranks_stay_identical(tmp_path, fast):
    from hetseq_9cme_b200.data import synthetic
    from hetseq_9cme_b200.distributed_utils import find_free_port
    d = str(tmp_path)
    synthetic.make_mnist(d, 16 * 7, 32)         # 7 batches over 2 ranks -> rank 1 gets one dummy (lock-step) batch
    port = find_free_port()
    mp.spawn(_worker, args=(2, port, d, d, fast), nprocs=2, join=True)
    a, b = torch.load(os.path.join(d, "rank0.pt")), torch.load(os.path.join(d, "rank1.pt"))
    for k in a:
        assert torch.equal(a[k], b[k]), k                          # replicas bit-identical after training
    m0, m1 = torch.load(os.path.join(d, "meta0.pt")), torch.load(os.path.join(d, "meta1.pt"))
    assert m0["updates"] == m1["updates"] == 4                     # ceil(7 / 2) steps on both ranks


def _single(data_dir, out_dir):
    sys.path.insert(0, ROOT)
    from hetseq_9cme_b200 import options, train
    argv = ["--task", "mnist", "--optimizer", "adadelta", "--data", data_dir, "--cpu", "--max-sentences", "16", "--lr", "1.0",
            "--valid-subset", "test", "--save-dir", os.path.join(out_dir, "ckpt1"), "--distributed-world-size", "1",
            "--num-workers", "0", "--clip-norm", "100", "--log-format", "none", "--max-epoch", "1", "--no-save", "--seed", "3",
            "--update-freq", "2", "--fast-stat-sync"]
    c = train.main(options.parse_training_args(argv))
    return {k: v.clone() for k, v in c.get_model().state_dict().items()}


def test_two_ranks_equal_one_rank_with_accumulation(tmp_path):
    """W=2 with batch b and W=1 with --update-freq 2 see the same samples per update (batches r, r+W), so with
    dropout disabled by eval-free determinism... the summed gradients, hence the weights, must agree closely."""
    from hetseq_9cme_b200.data import synthetic
    from hetseq_9cme_b200.distributed_utils import find_free_port
    import hetseq_9cme_b200.tasks.tasks as T
    d = str(tmp_path)
    synthetic.make_mnist(d, 16 * 8, 32)
    port = find_free_port()
    os.environ["B200SEQ_TEST_NO_DROPOUT"] = "1"
    try:
        mp.spawn(_worker, args=(2, port, d, d, True), nprocs=2, join=True)
        two = torch.load(os.path.join(d, "rank0.pt"))
        one = _single(d, d)
    finally:
        os.environ.pop("B200SEQ_TEST_NO_DROPOUT", None)
    for k in two:
        assert torch.allclose(two[k], one[k], atol=2e-5), k


def _comm_selftest_worker(rank, world, tag, out_dir):
    import ctypes
    from hetseq_9cme_b200.ops import native
    lib = native.cuda_lib()
    lib.b200_comm_selftest.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_char_p, ctypes.c_int]
    rc = lib.b200_comm_selftest(rank, world, tag.encode(), 20000)
    msg = lib.b200_last_error()
    with open(os.path.join(out_dir, f"rc{rank}"), "w") as f:
        f.write(f"{rc} {msg.decode() if rc and msg else ''}")


@pytest.mark.parametrize("world", [1, 3])
def test_comm_runtime_control_plane(tmp_path, world):
    """The runtime's host side without a GPU (csrc/comm_runtime.cu: b200_comm_selftest): rendezvous over abstract
    Unix sockets, AND-agreement, descriptors passed as SCM_RIGHTS and read back, rounds arriving out of order.
    The same code carries the CUDA VMM handles on a GPU box (tests/test_gpu_multi.py)."""
    from hetseq_9cme_b200.ops import native
    if not native.cuda_lib_available():
        pytest.skip("libb200seq.so not built")
    tag = f"selftest.{os.getpid()}.{world}"
    d = str(tmp_path)
    if world == 1:
        _comm_selftest_worker(0, 1, tag, d)
    else:
        mp.spawn(_comm_selftest_worker, args=(world, tag, d), nprocs=world, join=True)
    for r in range(world):
        assert open(os.path.join(d, f"rc{r}")).read().split(" ", 1)[0] == "0", open(os.path.join(d, f"rc{r}")).read()


def test_comm_runtime_reports_a_missing_rank(tmp_path):
    """A rank that never shows up is an error after the timeout, not a hang (csrc/comm_runtime.cu: send_msg)."""
    import ctypes
    from hetseq_9cme_b200.ops import native
    if not native.cuda_lib_available():
        pytest.skip("libb200seq.so not built")
    lib = native.cuda_lib()
    lib.b200_comm_selftest.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_char_p, ctypes.c_int]
    rc = lib.b200_comm_selftest(0, 2, f"selftest.alone.{os.getpid()}".encode(), 300)
    assert rc != 0 and b"cannot reach rank 1" in lib.b200_last_error()


def test_dead_rank_is_restarted_from_the_last_checkpoint(tmp_path):
    """--max-restarts: rank 1 is killed at update 5 (--fault-inject); the launcher stops rank 0, starts both again, they
    resume from the checkpoint written at update 4 and finish the epoch with the same number of updates as a run
    without the fault (recovery path the reference does not have: hetseq/train.py:196-250 just spawns)."""
    import subprocess
    from hetseq_9cme_b200.data import synthetic
    d = str(tmp_path)
    synthetic.make_mnist(d, 16 * 16, 32)        # 16 batches over 2 ranks -> 8 updates
    save = os.path.join(d, "ckpt")
    cmd = [sys.executable, os.path.join(ROOT, "hetseq_9cme_b200", "train.py"), "--task", "mnist", "--optimizer", "adadelta",
           "--data", d, "--cpu", "--distributed-backend", "gloo", "--distributed-world-size", "2", "--max-sentences", "16",
           "--lr", "1.0", "--valid-subset", "test", "--save-dir", save, "--num-workers", "0", "--clip-norm", "100",
           "--log-format", "simple", "--log-interval", "1", "--max-epoch", "1", "--seed", "3", "--fast-stat-sync",
           "--save-interval-updates", "2", "--fault-inject", "1:5", "--max-restarts", "1"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    out = r.stdout + r.stderr
    assert r.returncode == 0, out[-3000:]
    assert "fault injection: rank 1 exits at update 5" in out
    assert "restart 1/1 from the last checkpoint" in out
    assert "loaded checkpoint" in out and "@ 4 updates" in out          # resumed from the mid-epoch checkpoint
    last = torch.load(os.path.join(save, "checkpoint_last.pt"), map_location="cpu", weights_only=False)
    assert last["optimizer_history"][-1]["num_updates"] == 8


def test_heterogeneous_nodes_one_command_per_node(tmp_path):
    """The reference's launch model for nodes with different GPU counts (docs/source/distribute.rst, the `*_het` job
    scripts): one command per node, each with its own --distributed-gpus and the rank of its first process.  Here
    "node 0" brings two ranks and "node 1" one, on CPUs with gloo; all three meet at the same tcp:// init method."""
    import subprocess
    from hetseq_9cme_b200.data import synthetic
    from hetseq_9cme_b200.distributed_utils import find_free_port
    d = str(tmp_path)
    synthetic.make_mnist(d, 16 * 6, 32)                          # 6 batches over 3 ranks -> 2 updates
    port = find_free_port()
    common = [sys.executable, os.path.join(ROOT, "hetseq_9cme_b200", "train.py"), "--task", "mnist", "--optimizer", "adadelta",
              "--data", d, "--cpu", "--distributed-backend", "gloo", "--distributed-init-method", "tcp://127.0.0.1:%d" % port,
              "--distributed-world-size", "3", "--max-sentences", "16", "--lr", "1.0", "--valid-subset", "test", "--num-workers", "0",
              "--clip-norm", "100", "--log-format", "none", "--max-epoch", "1", "--seed", "3"]
    node0 = subprocess.Popen(common + ["--distributed-gpus", "2", "--distributed-rank", "0", "--save-dir", os.path.join(d, "ck0")],
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=ROOT)
    node1 = subprocess.Popen(common + ["--distributed-gpus", "1", "--distributed-rank", "2", "--save-dir", os.path.join(d, "ck1")],
                             stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, cwd=ROOT)
    try:
        out0, _ = node0.communicate(timeout=300)
        out1, _ = node1.communicate(timeout=300)
    finally:
        for p in (node0, node1):
            if p.poll() is None:
                p.kill()
    assert node0.returncode == 0 and node1.returncode == 0, out0[-2000:] + out1[-2000:]
    assert "initialized host" in out1 and "as rank 2" in out1
    last = torch.load(os.path.join(d, "ck0", "checkpoint_last.pt"), map_location="cpu", weights_only=False)
    assert last["optimizer_history"][-1]["num_updates"] == 2 and last["args"].distributed_world_size == 3
    assert not os.path.exists(os.path.join(d, "ck1", "checkpoint_last.pt"))      # only the master rank writes
