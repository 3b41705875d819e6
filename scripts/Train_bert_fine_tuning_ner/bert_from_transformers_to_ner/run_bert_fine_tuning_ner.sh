#!/bin/bash
# BERT token classification (NER) on CoNLL-2003-shaped files, warm-started from a HuggingFace transformers BERT state dict.
DIST=${DIST:-$(cd "$(dirname "$0")/../../.." && pwd)}
python3 ${DIST}/hetseq_9cme_b200/train.py \
  --task BertForTokenClassification --optimizer adam --lr-scheduler PolynomialDecayScheduler \
  --data ${DATA:-$DIST/ner_data} --dict ${DATA:-$DIST/ner_data}/vocab.txt --config_file ${CONFIG:-$DIST/bert_config.json} \
  --train_file ${DATA:-$DIST/ner_data}/train.txt --validation_file ${DATA:-$DIST/ner_data}/valid.txt \
  --test_file ${DATA:-$DIST/ner_data}/test.txt --max-sentences 32 --fast-stat-sync --max-epoch 4 --update-freq 1 \
  --valid-subset test --num-workers 4 --warmup-updates 100 --total-num-update 2000 --lr 5e-5 --weight-decay 0.01 \
  --pad-to-multiple-of 8 --transformers_state_dict ${CKPT:?set CKPT=pytorch_model.bin} \
  --distributed-world-size ${GPUS:-1} --save-dir bert_from_transformers_to_ner 2>&1 | tee bert_from_transformers_to_ner.log
