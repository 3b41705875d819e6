#!/bin/bash
# BERT phase-1 on one 8xB200 node: no --distributed-init-method -> the launcher picks a free local port and spawns
# one process per GPU; gradients are reduced by the NVLink peer-to-peer kernels.
DIST=${DIST:-$(cd "$(dirname "$0")/../../.." && pwd)}
DATA=${DATA:-$DIST/preprocessing/test_128/}
BERT=${BERT:-$DIST/preprocessing/uncased_L-12_H-768_A-12}
python3 ${DIST}/hetseq_9cme_b200/train.py \
  --task bert --data ${DATA} --dict ${BERT}/vocab.txt --config_file ${BERT}/bert_config.json \
  --max-sentences 32 --fast-stat-sync --max-update 450000 --update-freq 1 \
  --valid-subset test --num-workers 4 \
  --warmup-updates 10000 --total-num-update 1000000 --lr 0.0001 --weight-decay 0.01 \
  --distributed-world-size 8 --cuda-graph --max-predictions-per-seq 20 \
  --save-dir bert_node1gpu8 2>&1 | tee bert_node1gpu8.log
