#!/bin/bash
# MNIST example on 4 GPUs of one node (counterpart of STORE_RUN_FILE/Train_mnist/node1gpu4/main.sh).
DIST=${DIST:-$(cd "$(dirname "$0")/../../.." && pwd)}
python3 ${DIST}/hetseq_9cme_b200/train.py \
  --task mnist --optimizer adadelta --lr-scheduler PolynomialDecayScheduler \
  --data ${DATA:-$DIST/mnist_data} --clip-norm 100 \
  --max-sentences 64 --fast-stat-sync --max-epoch 20 --update-freq 1 \
  --valid-subset test --num-workers 4 \
  --warmup-updates 0 --total-num-update 50000 --lr 1.01 \
  --distributed-world-size 4 --save-dir mnist_node1gpu4 2>&1 | tee mnist_node1gpu4.log
