#!/bin/bash
# A/B of one environment switch on the headline bench (1 GPU): tools/gpu_call_ab.sh VAR valueA valueB [precision]
VAR=$1; A=$2; B=$3; PREC=${4:-tf32}
mkdir -p gpurun_out
timeout 900 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_ab.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_ab.log
for v in $A $B $A $B; do
  env $VAR=$v timeout 300 python bench.py --steps 30 --warmup 5 --precision $PREC --no-e2e > gpurun_out/ab_${VAR}_${v}.log 2>&1
  echo "$VAR=$v: $(grep -E '^\{' gpurun_out/ab_${VAR}_${v}.log | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.readline()); print(d['ms_per_step'], d['value'])")"
done
