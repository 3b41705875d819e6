#!/bin/bash
# data-path choice at 4 GPUs: BERT-large bf16 and BERT-base TF32 through the switch vs the peer forms
N=4; TAG=${1:-p4}
export B200_P2P_TIMEOUT_S=60 B200_COMM_TIMEOUT_S=60
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517"
run() { name=$1; shift; envs=$1; shift; env $envs timeout -k 10 240 $TR bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench_${name}_${N}gpu_${TAG}.log 2>&1; echo "$name rc=$?"; grep -E '^\{' gpurun_out/bench_${name}_${N}gpu_${TAG}.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print('   ', d['dtype'], d['value'], d['ms_per_step'], d.get('exposed_allreduce_ms_per_step'))"; }
run c4_nvls "B200_RS_PEER=0 B200_BCAST_PEER=0" --model bert-large --precision bf16
run c4_rsnvls_bcastpeer "B200_RS_PEER=0 B200_BCAST_PEER=1" --model bert-large --precision bf16
run c2_nvls "B200_RS_PEER=0 B200_BCAST_PEER=0"
