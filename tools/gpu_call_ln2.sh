#!/bin/bash
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "layernorm" > gpurun_out/ln_pytest2.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/ln_pytest2.log
for rep in a b; do
for mode in 0 2; do
  B200_LN_PIPE=$mode timeout 200 python bench.py --gpus 1 --steps 40 --warmup 5 > gpurun_out/bench_ln2_pipe${mode}_${rep}.log 2>gpurun_out/bench_ln2_pipe${mode}_${rep}.err
  echo "pipe=$mode rep=$rep rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_ln2_pipe${mode}_${rep}.log | head -1)"
done
done
