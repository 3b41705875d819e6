#!/bin/bash
# BERT-base NER fine-tuning on CoNLL-2003-style files, warm-started from a pre-training checkpoint (counterpart of
# STORE_RUN_FILE/Train_bert_fine_tuning_ner/bert_from_reddit_to_aida_ner/run_bert_fine_tuning_ner.sh).
DIST=${DIST:-$(cd "$(dirname "$0")/../.." && pwd)}
BERT=${BERT:-$DIST/preprocessing/uncased_L-12_H-768_A-12}
NER=${NER:-$DIST/preprocessing/conll2003}
python3 ${DIST}/hetseq_9cme_b200/train.py \
  --task BertForTokenClassification --optimizer adam --lr-scheduler PolynomialDecayScheduler \
  --data ${NER} --dict ${BERT}/vocab.txt --config_file ${BERT}/bert_config.json \
  --train_file ${NER}/train.txt --validation_file ${NER}/valid.txt --test_file ${NER}/test.txt \
  --hetseq_state_dict ${CKPT:-bert_node1gpu8/checkpoint_last.pt} --load_state_dict_strict False \
  --max-sentences 32 --fast-stat-sync --max-epoch 4 --update-freq 1 --valid-subset test --num-workers 4 \
  --warmup-updates 0 --total-num-update 100000 --lr 0.0001 --weight-decay 0.01 --find-unused-parameters \
  --distributed-world-size 1 --save-dir bert_ner 2>&1 | tee bert_ner.log
python3 ${DIST}/hetseq_9cme_b200/eval_bert_fine_tuning_ner.py --model_ckpt bert_ner/checkpoint_last.pt \
  --config_file ${BERT}/bert_config.json --dict ${BERT}/vocab.txt --test_file ${NER}/test.txt
