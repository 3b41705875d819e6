#!/usr/bin/env python3
"""Evaluate a token-classification checkpoint on a CoNLL-style test file: accuracy, precision, recall, F1 at the
entity level.  Working counterpart of the reference's stub hetseq/eval_bert_fine_tuning_ner.py (13 lines, raises
NameError) and of its hand-run script test/test_eval_bert_fine_tuning.py:39-169."""
from __future__ import annotations

import argparse
import os
import sys

if __name__ == "__main__" and __package__ is None:
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    __package__ = "hetseq_9cme_b200"

import torch

from . import seqeval_lite
from .bert_modeling import BertConfig, BertForTokenClassification
from .data import conll
from .data_collator import YD_DataCollatorForTokenClassification


def evaluate(model_ckpt, config_file, vocab, test_file, label_list=None, batch_size=32, device=None):
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    state = torch.load(model_ckpt, map_location="cpu", weights_only=False)
    sd = state["model"] if "model" in state else state
    sents = conll.read_conll(test_file)
    labels = label_list or getattr(state.get("args", None), "label_list", None) or conll.label_list_from(sents)
    model = BertForTokenClassification(BertConfig.from_json_file(config_file), len(labels))
    model.load_state_dict(sd, strict=True)
    model.to(device).eval()
    tok = conll.build_tokenizer(vocab)
    feats = conll.tokenize_and_align(tok, sents, {l: i for i, l in enumerate(labels)})
    coll = YD_DataCollatorForTokenClassification(tok)
    y_true, y_pred = [], []
    with torch.no_grad():
        for s in range(0, len(feats), batch_size):
            b = coll(feats[s:s + batch_size])
            lab = b.pop("labels")
            logits = model(**{k: v.to(device) for k, v in b.items()})
            pred = logits.argmax(-1).cpu()
            for p_row, l_row in zip(pred.tolist(), lab.tolist()):
                keep = [(labels[p], labels[l]) for p, l in zip(p_row, l_row) if l != -100]
                if keep:
                    y_pred.append([k[0] for k in keep])
                    y_true.append([k[1] for k in keep])
    out = {"accuracy_score": seqeval_lite.accuracy_score(y_true, y_pred), "precision": seqeval_lite.precision_score(y_true, y_pred),
           "recall": seqeval_lite.recall_score(y_true, y_pred), "f1": seqeval_lite.f1_score(y_true, y_pred)}
    print(out)
    return out


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--model_ckpt", required=True)
    ap.add_argument("--config_file", required=True)
    ap.add_argument("--dict", required=True)
    ap.add_argument("--test_file", required=True)
    ap.add_argument("--batch_size", type=int, default=32)
    a = ap.parse_args(argv)
    evaluate(a.model_ckpt, a.config_file, a.dict, a.test_file, batch_size=a.batch_size)


if __name__ == "__main__":
    main()
