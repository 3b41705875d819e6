#!/bin/bash
# Two nodes x 8 GPUs, node 0 (counterpart of STORE_RUN_FILE/Train_bert/node2gpu4/node2gpu4_main.sh).  Run sub.sh on
# the second node.  Across nodes the gradient reduction falls back to torch.distributed/NCCL (peer memory is a
# single-NVSwitch-domain feature).
DIST=${DIST:-$(cd "$(dirname "$0")/../../.." && pwd)}
DATA=${DATA:-$DIST/preprocessing/test_128/}
BERT=${BERT:-$DIST/preprocessing/uncased_L-12_H-768_A-12}
MASTER=${MASTER:-tcp://10.0.0.1:12345}
python3 ${DIST}/hetseq_9cme_b200/train.py \
  --task bert --data ${DATA} --dict ${BERT}/vocab.txt --config_file ${BERT}/bert_config.json \
  --max-sentences 32 --fast-stat-sync --max-update 450000 --update-freq 4 \
  --valid-subset test --num-workers 4 \
  --warmup-updates 10000 --total-num-update 1000000 --lr 0.0001 --weight-decay 0.01 \
  --distributed-init-method ${MASTER} --distributed-world-size 16 --distributed-gpus 8 --distributed-rank 0 \
  --save-dir bert_node2gpu8 2>&1 | tee bert_node2gpu8_main.log
