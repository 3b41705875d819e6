#!/bin/bash
# BASELINE configs 3-5 at N = 2 or 4 GPUs (ours), completing the 1/2/4/8 matrix
N=$1; TAG=${2:-mm}
export B200_P2P_TIMEOUT_S=60 B200_COMM_TIMEOUT_S=60
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29516"
run() { name=$1; shift; timeout -k 10 240 $TR bench.py --gpus $N --steps 20 --warmup 5 "$@" > gpurun_out/bench_${name}_${N}gpu_${TAG}.log 2>&1; echo "$name rc=$?"; grep -E '^\{' gpurun_out/bench_${name}_${N}gpu_${TAG}.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print('   ', d['dtype'], d['value'], d['ms_per_step'], d.get('replicas_identical'))"; }
run c3_ours_tf32 --seq-len 512 --batch 8
run c4_ours_bf16 --model bert-large --precision bf16
run c5_ours_tf32 --task ner
