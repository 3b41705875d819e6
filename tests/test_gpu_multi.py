"""Multi-GPU tests (skipped on boxes with one GPU): the communication runtime and the peer-to-peer kernels against
NCCL, data-parallel training through the Controller in every reducer mode against a single-process run on the same
global batches, CUDA-graph replay next to eagerly stepping ranks, sharded checkpoints, and the watchdog that turns a
dead peer into an error (SURVEY §4 "GPU comm tests", §5.2, §5.3)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
# small buckets: BERT-tiny then spans ~10 buckets and the tied word-embedding / decoder weight sits alone in one
SMALL_BUCKETS = {"B200_BUCKET_BYTES": str(256 * 1024)}


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _torchrun(n, script_args, env=None, timeout=420):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + os.getpid() % 200)] + script_args
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run(cmd, cwd=ROOT, env=e, capture_output=True, text=True, timeout=timeout)


def _ok(r, marker):
    assert r.returncode == 0 and marker in r.stdout, r.stdout[-3000:] + r.stderr[-3000:]


needs2 = pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")


@needs2
def test_p2p_kernels_match_nccl_and_replicated_adam():
    """All-reduce mode vs ncclAllReduce; reduce-scatter + sharded Adam + parameter broadcast vs all-reduce + the
    replicated Adam kernel, fp32 masters and bf16 working copies; optimizer state gather."""
    _ok(_torchrun(min(_ngpu(), 8), ["tools/gpu_probe_p2p.py", "--quick"]), "P2P PROBE PASS")


@needs2
def test_sharded_eager_steps_verified_bucket_by_bucket(tmp_path):
    """Eager steps with B200_P2P_DEBUG=1 (every bucket's slice vs the NCCL sum, epochs vs the host's count), compared
    with a single-process run on the same global batches, then a checkpoint round trip of the sharded state."""
    env = dict(SMALL_BUCKETS, B200_P2P_DEBUG="1")
    _ok(_torchrun(2, ["tools/dp_debug_run.py", str(tmp_path), "--mode", "sharded", "--compare-single", "--checkpoint"], env=env),
        "DP DEBUG RUN PASS")


@needs2
def test_graph_replay_next_to_eager_ranks_with_accumulation(tmp_path):
    """--cuda-graph --update-freq 2 on uneven shards: at the end of each epoch one rank gets the empty filler batch and
    steps eagerly while the other replays or captures — both must issue the same reductions."""
    _ok(_torchrun(2, ["tools/dp_debug_run.py", str(tmp_path), "--mode", "sharded", "--graph", "--update-freq", "2", "--uneven",
                      "--steps", "10", "--compare-single"], env=SMALL_BUCKETS), "DP DEBUG RUN PASS")


@needs2
def test_replicated_all_reduce_mode(tmp_path):
    _ok(_torchrun(2, ["tools/dp_debug_run.py", str(tmp_path), "--mode", "replicated", "--graph", "--compare-single"],
                  env=SMALL_BUCKETS), "DP DEBUG RUN PASS")


@needs2
def test_bf16_sharded_masters_and_checkpoint(tmp_path):
    _ok(_torchrun(2, ["tools/dp_debug_run.py", str(tmp_path), "--mode", "sharded", "--graph", "--precision", "bf16",
                      "--compare-single", "--checkpoint"], env=SMALL_BUCKETS), "DP DEBUG RUN PASS")


@needs2
def test_nccl_reducer_path(tmp_path):
    """--comm torch: NCCL all-reduces per bucket, replicated optimizer (the portable baseline of the p2p path)."""
    _ok(_torchrun(2, ["tools/dp_debug_run.py", str(tmp_path), "--mode", "torch", "--compare-single"]), "DP DEBUG RUN PASS")


@needs2
def test_pipelined_update_matches(tmp_path):
    """--pipelined-update (here through its environment switch): the sharded update parked at the end of the step and
    issued per bucket under the next forward pass; uneven shards make ranks flush and step eagerly in between."""
    _ok(_torchrun(2, ["tools/dp_debug_run.py", str(tmp_path), "--mode", "sharded", "--graph", "--uneven", "--steps", "10",
                      "--compare-single", "--checkpoint"], env=dict(SMALL_BUCKETS, B200_PIPELINED_UPDATE="1")), "DP DEBUG RUN PASS")


@needs2
def test_dead_peer_becomes_an_error_not_a_hang(tmp_path):
    """Rank 1 exits before its 4th step.  Rank 0 is an independent process (no launcher that would kill it): its
    reduction kernel waits B200_P2P_TIMEOUT_S for the peer, records who is missing, traps; the process reports it."""
    port = str(29850 + os.getpid() % 100)
    procs = []
    for rank in range(2):
        e = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1", MASTER_PORT=port,
                 B200_P2P_TIMEOUT_S="4")
        procs.append(subprocess.Popen([sys.executable, "tools/dp_debug_run.py", str(tmp_path), "--mode", "sharded", "--steps", "8",
                                       "--fault-inject", "1:3"], cwd=ROOT, env=e, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        try:
            outs.append(p.communicate(timeout=240)[0])
        except subprocess.TimeoutExpired:
            p.kill()
            outs.append(p.communicate()[0] + "\n<<killed by the test: still running after 240 s>>")
    assert procs[1].returncode == 17, outs[1][-1500:]
    assert procs[0].returncode == 86, outs[0][-3000:]
    assert "WATCHDOG REPORT: rank 1 did not reach" in outs[0], outs[0][-3000:]
