#!/bin/bash
# BERT-large bf16 on 2 GPUs: optimizer placement and reduction grid
N=2; TAG=${1:-l2}
export B200_P2P_TIMEOUT_S=60 B200_COMM_TIMEOUT_S=60
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29518"
run() { name=$1; shift; envs=$1; shift; env $envs timeout -k 10 240 $TR bench.py --gpus $N --steps 20 --warmup 5 --model bert-large --precision bf16 "$@" > gpurun_out/bench_c4_${name}_${N}gpu_${TAG}.log 2>&1; echo "$name rc=$?"; grep -E '^\{' gpurun_out/bench_c4_${name}_${N}gpu_${TAG}.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print('   ', d['dtype'], d['value'], d['ms_per_step'], d.get('exposed_allreduce_ms_per_step'))"; }
run replicated "B200_NO_SHARDED=1"
run sharded_blocks96 "B200_P2P_BLOCKS=96"
run sharded_noverlap "B200_P2P_NO_OVERLAP=1"
