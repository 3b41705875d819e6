#!/bin/bash
# One multi-GPU measurement call: runtime + kernels probe, multi-GPU tests, bench at N GPUs and at 1 GPU.
# usage: tools/gpu_call_multi.sh NGPU [tag]
N=${1:-2}; TAG=${2:-r2}
export B200_P2P_TIMEOUT_S=${B200_P2P_TIMEOUT_S:-30} B200_COMM_TIMEOUT_S=60
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
nvidia-smi --query-gpu=index,name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/smi_${TAG}.csv 2>&1
timeout -k 10 400 $TR --master-port 29511 tools/gpu_probe_p2p.py > gpurun_out/probe_${N}gpu_${TAG}.log 2>&1; echo "probe rc=$?"
tail -25 gpurun_out/probe_${N}gpu_${TAG}.log
timeout -k 10 1200 python -m pytest tests/test_gpu_multi.py -q -m gpu > gpurun_out/pytest_multi_${TAG}.log 2>&1; echo "pytest multi rc=$?"
tail -40 gpurun_out/pytest_multi_${TAG}.log
timeout -k 10 300 $TR --master-port 29512 bench.py --gpus $N --steps 20 --warmup 5 > gpurun_out/bench_${N}gpu_${TAG}.log 2>&1; echo "bench N rc=$?"
tail -3 gpurun_out/bench_${N}gpu_${TAG}.log
