"""Multi-GPU tests (skipped on boxes with one GPU): the peer-to-peer reducer against NCCL, and a data-parallel
training run with the reducer's debug verification switched on (SURVEY §4 "GPU comm tests", §5.2)."""
import os
import subprocess
import sys

import pytest
import torch

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _ngpu():
    return torch.cuda.device_count() if torch.cuda.is_available() else 0


def _torchrun(n, script_args, env=None, timeout=300):
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(n), "--master-addr", "127.0.0.1",
           "--master-port", str(29600 + os.getpid() % 200)] + script_args
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run(cmd, cwd=ROOT, env=e, capture_output=True, text=True, timeout=timeout)


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
def test_p2p_reducer_matches_nccl():
    n = min(_ngpu(), 8)
    r = _torchrun(n, ["tools/gpu_probe_p2p.py"])
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "P2P PROBE PASS" in r.stdout


@pytest.mark.skipif(_ngpu() < 2, reason="needs at least 2 GPUs")
def test_data_parallel_training_with_reducer_verification(tmp_path):
    """Eight eager steps of BERT-tiny on 2 GPUs with B200_P2P_DEBUG=1: every bucket of every step is compared with the
    NCCL sum of its pre-launch copy, replicas must be bit-identical and epochs must match the host's count."""
    r = _torchrun(2, ["tools/dp_debug_run.py", str(tmp_path)], env={"B200_P2P_DEBUG": "1"})
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "DP DEBUG RUN PASS" in r.stdout
