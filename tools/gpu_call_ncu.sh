#!/bin/bash
# `ncu --set full` of one or two steady-state launches of each non-GEMM kernel of the training step (one GPU).
TAG=${1:-n1}
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-e2e"
$CMD > gpurun_out/plain_ncu_${TAG}.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/plain_ncu_${TAG}.log; exit 1; }
for k in ln_fwd_kernel adam_flat_kernel sumsq_kernel ce_fwd_kernel ce_bwd_kernel embed_scatter_kernel embed_sum_kernel attn_fwd_wide_kernel attn_bwd_block_kernel; do
  ncu --set full --clock-control none --import-source on -k regex:$k -s 2 -c 2 -o gpurun_out/prof_${k}_${TAG} $CMD > gpurun_out/ncu_${k}_${TAG}.log 2>&1
  echo "ncu $k rc=$?"
done
