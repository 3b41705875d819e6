#!/bin/bash
# One multi-GPU measurement call: runtime + kernels probe (with the data-path sweep), multi-GPU tests, bench at N GPUs
# in the three optimizer placements.   usage: tools/gpu_call_multi.sh NGPU [tag] [notests]
N=${1:-2}; TAG=${2:-r2}; NOTESTS=${3:-}
export B200_P2P_TIMEOUT_S=${B200_P2P_TIMEOUT_S:-30} B200_COMM_TIMEOUT_S=60
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout -k 10 400 $TR --master-port 29511 tools/gpu_probe_p2p.py > gpurun_out/probe_${N}gpu_${TAG}.log 2>&1; echo "probe rc=$?"
grep -E "PROBE|BW|SWEEP|sharded|allreduce step" gpurun_out/probe_${N}gpu_${TAG}.log | tail -150
if [ -z "$NOTESTS" ]; then
  timeout -k 10 1500 python -m pytest tests/test_gpu_multi.py -q -m gpu > gpurun_out/pytest_multi_${TAG}.log 2>&1; echo "pytest multi rc=$?"
  tail -30 gpurun_out/pytest_multi_${TAG}.log
fi
for mode in pipelined nodefer replicated; do
  case $mode in
    pipelined) ENVS="B200_PIPELINED_UPDATE=1";;
    nodefer) ENVS="";;
    replicated) ENVS="B200_NO_SHARDED=1";;
  esac
  env $ENVS timeout -k 10 300 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_${N}gpu_${mode}_${TAG}.log 2>&1; echo "bench $mode rc=$?"
  tail -1 gpurun_out/bench_${N}gpu_${mode}_${TAG}.log | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print({k:d[k] for k in ('value','ms_per_step','final_loss','exposed_allreduce_ms_per_step','replicas_identical') if k in d}, d.get('e2e',{}).get('value'))" || tail -5 gpurun_out/bench_${N}gpu_${mode}_${TAG}.log
done
