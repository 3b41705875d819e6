#!/bin/bash
# 1-GPU rows of configs 3 and 5 on the final build.
mkdir -p gpurun_out
timeout 90 python bench.py --gpus 1 --steps 20 --warmup 5 --seq-len 512 --batch 8 > gpurun_out/bench_c3_final_c.log 2>gpurun_out/bench_c3_final_c.err
echo "c3 rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_c3_final_c.log | head -1)"
timeout 90 python bench.py --gpus 1 --steps 20 --warmup 5 --task ner > gpurun_out/bench_c5_final_c.log 2>gpurun_out/bench_c5_final_c.err
echo "c5 rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_c5_final_c.log | head -1)"
