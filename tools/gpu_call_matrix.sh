#!/bin/bash
# BASELINE.md section B matrix at N GPUs: every BASELINE.json config, ours and (where the reference can run it) the
# unmodified reference.   usage: tools/gpu_call_matrix.sh NGPU tag [noref]
N=${1:-1}; TAG=${2:-m}; NOREF=${3:-}
export B200_P2P_TIMEOUT_S=${B200_P2P_TIMEOUT_S:-60} B200_COMM_TIMEOUT_S=60
mkdir -p gpurun_out
if [ "$N" -gt 1 ]; then RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513"; else RUN="python"; fi
summ() { tail -1 "$1" | python -c "
import sys, json
try:
    d = json.loads(sys.stdin.readline())
    print('   ', d.get('impl'), d.get('dtype'), {k: d[k] for k in ('value', 'ms_per_step', 'final_loss', 'replicas_identical', 'unavailable') if k in d}, 'e2e', d.get('e2e', {}).get('value'), 'ref_tf32', (d.get('reference_tf32') or {}).get('value'))
except Exception as e:
    print('    no JSON line:', e)
" ; }
one() {  # name, args...
  name=$1; shift
  timeout -k 10 ${RUN_TIMEOUT:-600} $RUN bench.py --gpus $N --steps ${STEPS:-20} --warmup 5 "$@" > gpurun_out/bench_${name}_${N}gpu_${TAG}.log 2>&1; echo "$name rc=$?"
  summ gpurun_out/bench_${name}_${N}gpu_${TAG}.log
}
if [ "$N" -gt 1 ]; then
  # multi-GPU: the headline precision of each config (bf16 for BERT-large, as BASELINE.json names it); NER last (its
  # ranks capture different batch shapes)
  one c2_ours_tf32
  one c3_ours_tf32 --seq-len 512 --batch 8
  one c4_ours_bf16 --model bert-large --precision bf16
  if [ -z "$NOREF" ]; then
    one c2_ref --impl reference
    one c3_ref --impl reference --seq-len 512 --batch 8
    one c4_ref --impl reference --model bert-large
    one c5_ref --impl reference --task ner
  fi
  one c5_ours_tf32 --task ner
else
  one c2_ours_tf32
  one c2_ours_bf16 --precision bf16
  one c3_ours_tf32 --seq-len 512 --batch 8
  one c3_ours_bf16 --seq-len 512 --batch 8 --precision bf16
  one c4_ours_bf16 --model bert-large --precision bf16
  one c4_ours_tf32 --model bert-large
  one c5_ours_tf32 --task ner
  if [ -z "$NOREF" ]; then
    one c2_ref --impl reference
    one c3_ref --impl reference --seq-len 512 --batch 8
    one c4_ref --impl reference --model bert-large
    one c5_ref --impl reference --task ner
  fi
fi
