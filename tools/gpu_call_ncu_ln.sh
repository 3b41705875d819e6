#!/bin/bash
# ncu --set full of the streamed LayerNorm backward (two steady-state launches), after a plain run of the same command.
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-e2e"
timeout 60 $CMD > gpurun_out/plain_ncu_ln.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_ncu_ln.log; exit 1; }
echo "plain ok"
timeout 60 ncu --set full --clock-control none --import-source on -k regex:ln_bwd2_kernel -s 2 -c 2 -o gpurun_out/prof_ln_bwd2_pipe $CMD > gpurun_out/ncu_ln_bwd2_pipe.log 2>&1
echo "ncu rc=$?"
