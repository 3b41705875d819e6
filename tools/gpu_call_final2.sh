#!/bin/bash
# final check on 2 GPUs: the multi-GPU tests, the headline bench, BERT-large bf16 with the automatic reduction grid
export B200_P2P_TIMEOUT_S=60 B200_COMM_TIMEOUT_S=60
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29519"
timeout -k 10 900 python -m pytest tests/test_gpu_multi.py -q -m gpu > gpurun_out/pytest_multi_final.log 2>&1; echo "pytest multi rc=$?"; tail -3 gpurun_out/pytest_multi_final.log
for cfg in "c2_tf32" "c4_bf16 --model bert-large --precision bf16"; do
  set -- $cfg; name=$1; shift
  timeout -k 10 240 $TR bench.py --gpus 2 --steps 20 --warmup 5 "$@" > gpurun_out/bench_${name}_2gpu_final.log 2>&1; echo "$name rc=$?"
  grep -E '^\{' gpurun_out/bench_${name}_2gpu_final.log | tail -1 | python -c "
import sys,json
d=json.loads(sys.stdin.readline()); print('   ', d['dtype'], d['value'], d['ms_per_step'], d.get('exposed_allreduce_ms_per_step'), d.get('replicas_identical'))"
done
