#!/bin/bash
# BERT phase-1 pre-training on one GPU (counterpart of the reference's STORE_RUN_FILE/Train_bert/node1gpu1/main.sh).
DIST=${DIST:-$(cd "$(dirname "$0")/../../.." && pwd)}
DATA=${DATA:-$DIST/preprocessing/test_128/}
BERT=${BERT:-$DIST/preprocessing/uncased_L-12_H-768_A-12}
python3 ${DIST}/hetseq_9cme_b200/train.py \
  --task bert --data ${DATA} --dict ${BERT}/vocab.txt --config_file ${BERT}/bert_config.json \
  --max-sentences 32 --fast-stat-sync --max-update 900000 --update-freq 4 \
  --valid-subset test --num-workers 4 \
  --warmup-updates 10000 --total-num-update 1000000 --lr 0.0001 --weight-decay 0.01 \
  --distributed-world-size 1 --device-id 0 --cuda-graph --max-predictions-per-seq 20 \
  --save-dir bert_node1gpu1 2>&1 | tee bert_node1gpu1.log
