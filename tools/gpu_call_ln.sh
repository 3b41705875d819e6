#!/bin/bash
# One GPU call: correctness of the streaming LayerNorm kernels, their device time, and the step time with them.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_kernels.py -x -q -m gpu -k "layernorm" > gpurun_out/ln_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/ln_pytest.log
timeout 120 python tools/gpu_probe_ln.py > gpurun_out/ln_probe.log 2>&1; echo "probe rc=$?"
timeout 120 python tools/gpu_probe_ln.py --cols 1024 --dtype bf16 >> gpurun_out/ln_probe.log 2>&1
cat gpurun_out/ln_probe.log
for rep in a b; do
for mode in 0 3; do
  B200_LN_PIPE=$mode timeout 200 python bench.py --gpus 1 --steps 30 --warmup 5 > gpurun_out/bench_ln_pipe${mode}_${rep}.log 2>gpurun_out/bench_ln_pipe${mode}_${rep}.err
  echo "pipe=$mode rep=$rep rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_ln_pipe${mode}_${rep}.log | head -1)"
done
done
B200_LN_PIPE=3 B200_LN_PIPE_FWD_CTAS=1 B200_LN_PIPE_BWD_CTAS=1 timeout 200 python bench.py --gpus 1 --steps 30 --warmup 5 > gpurun_out/bench_ln_pipe3_c1.log 2>gpurun_out/bench_ln_pipe3_c1.err
echo "pipe=3 ctas=1 rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_ln_pipe3_c1.log | head -1)"
