#!/bin/bash
# bf16 step time with the CUDA library built from older commits (B200SEQ_LIB), to locate a regression.
mkdir -p gpurun_out
for c in head 860e7c5 6d1ae49 eebb807; do
  if [ $c = head ]; then unset B200SEQ_LIB; else export B200SEQ_LIB=$PWD/hetseq_9cme_b200/lib/variants/libb200seq_$c.so; fi
  timeout 200 python bench.py --gpus 1 --steps 20 --warmup 5 --precision bf16 > gpurun_out/bench_bisect_bf16_$c.log 2>gpurun_out/bench_bisect_bf16_$c.err
  echo "lib=$c rc=$? $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/bench_bisect_bf16_$c.log | head -1)"
done
