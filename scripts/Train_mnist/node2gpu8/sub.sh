#!/bin/bash
# MNIST on 2 nodes x 4 GPUs, node 0; sub.sh = same with --distributed-rank 4 (heterogeneous GPU types are fine: every
# rank runs the same program, the slowest rank sets the pace).
DIST=${DIST:-$(cd "$(dirname "$0")/../../.." && pwd)}
python3 ${DIST}/hetseq_9cme_b200/train.py \
  --task mnist --optimizer adadelta --data ${DATA:-$DIST/mnist_data} --clip-norm 100 \
  --max-sentences 64 --fast-stat-sync --max-epoch 20 --valid-subset test --num-workers 4 --lr 1.01 \
  --distributed-init-method ${MASTER:-tcp://10.0.0.1:11111} --distributed-world-size 8 --distributed-gpus 4 --distributed-rank 4 \
  --save-dir mnist_node2gpu8 2>&1 | tee mnist_node2gpu8_sub1.log
