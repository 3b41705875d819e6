#!/bin/bash
# what the driver runs at round end, on one GPU: the GPU test-suite, smoke(), bench.py in both arms
mkdir -p gpurun_out
timeout -k 10 1200 python -m pytest tests -x -q -m gpu > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?"; tail -2 gpurun_out/smoke_final.log
timeout 400 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ref_final.log 2>&1; echo "ref rc=$?"; grep -E '^\{' gpurun_out/bench_ref_final.log | tail -1 | cut -c1-300
timeout 300 python bench.py --gpus 1 --steps 20 --warmup 5 > gpurun_out/bench_ours_final.log 2>&1; echo "ours rc=$?"; grep -E '^\{' gpurun_out/bench_ours_final.log | tail -1 | cut -c1-300
