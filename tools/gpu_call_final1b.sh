#!/bin/bash
# Final 1-GPU validation of HEAD: GPU test-suite, smoke(), bench (TF32), bf16 with / without the streamed LN backward.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_final_b.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/pytest_gpu_final_b.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final_b.log 2>&1; echo "smoke rc=$?"; tail -1 gpurun_out/smoke_final_b.log
timeout 200 python bench.py --gpus 1 --steps 40 --warmup 5 > gpurun_out/bench_ours_final_b.log 2>gpurun_out/bench_ours_final_b.err; echo "ours rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_ours_final_b.log | head -1)"
for mode in 0 2; do
  B200_LN_PIPE=$mode timeout 200 python bench.py --gpus 1 --steps 40 --warmup 5 --precision bf16 > gpurun_out/bench_bf16_lnpipe${mode}.log 2>gpurun_out/bench_bf16_lnpipe${mode}.err
  echo "bf16 pipe=$mode rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_bf16_lnpipe${mode}.log | head -1)"
done
