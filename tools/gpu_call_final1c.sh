#!/bin/bash
# Validation after restoring the single-box TMA producer: GEMM / LN / attention kernel tests, then both precisions.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_final_c.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu_final_c.log
for prec in tf32 bf16; do
  timeout 200 python bench.py --gpus 1 --steps 30 --warmup 5 --precision $prec > gpurun_out/bench_final_c_$prec.log 2>gpurun_out/bench_final_c_$prec.err
  echo "$prec rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_final_c_$prec.log | head -1)"
done
timeout 200 python bench.py --gpus 1 --steps 20 --warmup 5 --precision bf16 --model bert-large > gpurun_out/bench_final_c_large.log 2>gpurun_out/bench_final_c_large.err
echo "large bf16 rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_final_c_large.log | head -1)"
