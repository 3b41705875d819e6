#!/bin/bash
# 8-GPU: reference NER arm, and the like-for-like comparison of the reducers with eagerly issued steps (NCCL cannot be
# captured with the step here, so both arms issue every kernel from Python).
N=8; TAG=${1:-f8}
export B200_P2P_TIMEOUT_S=60 B200_COMM_TIMEOUT_S=60
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514"
run() { name=$1; shift; timeout -k 10 300 $TR bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench_${name}_8gpu_${TAG}.log 2>&1; echo "$name rc=$?"; grep -E '^\{' gpurun_out/bench_${name}_8gpu_${TAG}.log | tail -1 | cut -c1-360; }
run c5_ref --impl reference --task ner
run c2_eager_p2p --no-graph
run c2_eager_nccl --no-graph --comm torch
