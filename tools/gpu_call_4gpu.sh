#!/bin/bash
# 4-GPU box: smoke() of the graft entry, the multi-GPU tests (probe runs on all 4), bench at 4 GPUs (both arms)
TAG=${1:-q4}
export B200_P2P_TIMEOUT_S=60 B200_COMM_TIMEOUT_S=60
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_${TAG}.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke_${TAG}.log
timeout -k 10 1200 python -m pytest tests/test_gpu_multi.py -q -m gpu > gpurun_out/pytest_multi_${TAG}.log 2>&1; echo "pytest multi rc=$?"; tail -4 gpurun_out/pytest_multi_${TAG}.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29515"
timeout -k 10 300 $TR bench.py --gpus 4 --steps 20 --warmup 5 > gpurun_out/bench_c2_ours_tf32_4gpu_${TAG}.log 2>&1; echo "bench rc=$?"
grep -E '^\{' gpurun_out/bench_c2_ours_tf32_4gpu_${TAG}.log | tail -1 | cut -c1-330
timeout -k 10 300 $TR bench.py --gpus 4 --steps 20 --warmup 5 --impl reference > gpurun_out/bench_c2_ref_4gpu_${TAG}.log 2>&1; echo "ref rc=$?"
grep -E '^\{' gpurun_out/bench_c2_ref_4gpu_${TAG}.log | tail -1 | cut -c1-330
