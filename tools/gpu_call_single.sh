#!/bin/bash
# One single-GPU call: the GPU test-suite, bench lines (tf32 / bf16), a launch list of one eagerly issued step and
# `ncu --set full` captures of the memory-bound kernels.   usage: tools/gpu_call_single.sh [tag] [skip-ncu]
TAG=${1:-g1}; SKIPNCU=${2:-}
mkdir -p gpurun_out
timeout -k 10 1500 python -m pytest tests -q -m gpu -x > gpurun_out/pytest_gpu_${TAG}.log 2>&1; echo "pytest rc=$?"
tail -15 gpurun_out/pytest_gpu_${TAG}.log
for prec in tf32 bf16; do
  timeout -k 10 300 python bench.py --steps 20 --warmup 5 --precision $prec > gpurun_out/bench_1gpu_${prec}_${TAG}.log 2>&1; echo "bench $prec rc=$?"
  tail -1 gpurun_out/bench_1gpu_${prec}_${TAG}.log | cut -c1-400
done
if [ -z "$SKIPNCU" ] || [ "$SKIPNCU" = "listonly" ]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-graph --no-e2e"
  $CMD > gpurun_out/plain_${TAG}.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -s 880 -c 620 --csv --log-file gpurun_out/launches_${TAG}.csv $CMD > gpurun_out/ncu_list_${TAG}.log 2>&1
  echo "ncu list rc=$?"
  [ "$SKIPNCU" = "listonly" ] || { $CMD > gpurun_out/plain2_${TAG}.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:'ln_fwd_kernel|ln_bwd2_kernel|adam_flat_kernel|ce_fwd_kernel|ce_bwd_kernel|embed_scatter_kernel|colsum_atomic_kernel' -s 30 -c 14 -o gpurun_out/prof_membound_${TAG} $CMD > gpurun_out/ncu_full_${TAG}.log 2>&1
  echo "ncu full rc=$?"; }
fi
# BASELINE config 5 on one GPU, both arms
timeout -k 10 400 python bench.py --steps 20 --warmup 5 --task ner > gpurun_out/bench_c5_ours_tf32_1gpu_${TAG}.log 2>&1; echo "ner ours rc=$?"
grep -E '^\{' gpurun_out/bench_c5_ours_tf32_1gpu_${TAG}.log | tail -1 | cut -c1-330
timeout -k 10 400 python bench.py --steps 20 --warmup 5 --task ner --impl reference > gpurun_out/bench_c5_ref_1gpu_${TAG}.log 2>&1; echo "ner ref rc=$?"
grep -E '^\{' gpurun_out/bench_c5_ref_1gpu_${TAG}.log | tail -1 | cut -c1-330
for v in B200_SIDE_COLSUM B200_LOCAL_NORM B200_GEMM_CTA2; do
  PREC=tf32; [ $v = B200_GEMM_CTA2 ] && PREC=bf16
  env $v=1 timeout -k 10 300 python bench.py --steps 20 --warmup 5 --precision $PREC > gpurun_out/bench_1gpu_${v}_${TAG}.log 2>&1
  echo "$v: $(grep -E '^\{' gpurun_out/bench_1gpu_${v}_${TAG}.log | tail -1 | cut -c1-200)"
done
