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


def to_transformers_state_dict(sd):
    """This framework's (= the reference's) BERT keys -> Hugging Face ``transformers`` keys."""
    out = {}
    for k, v in sd.items():
        k = k.replace("intermediate.dense_act.", "intermediate.dense.").replace("pooler.dense_act.", "pooler.dense.")
        k = k.replace("cls.predictions.transform.dense_act.", "cls.predictions.transform.dense.")
        out[k] = v
    return out


def _build_model(impl, config_file, num_labels, sd):
    """``impl='native'``: this framework's BertForTokenClassification.  ``impl='transformers'``: the Hugging Face model
    loaded with the same checkpoint — the cross-implementation check of the reference's
    test/TRANFORMERS_test_eval_bert_fine_tuning.py:1-288 (a HetSeq-trained checkpoint must evaluate identically under
    an independent BERT implementation)."""
    if impl == "native":
        model = BertForTokenClassification(BertConfig.from_json_file(config_file), num_labels)
        model.load_state_dict(sd, strict=True)
        return model, (lambda m, batch: m(**batch))
    import json
    from transformers import BertConfig as HFConfig, BertForTokenClassification as HFModel
    with open(config_file) as f:
        cfg = json.load(f)
    cfg["num_labels"] = num_labels
    cfg.setdefault("layer_norm_eps", 1e-12)
    model = HFModel(HFConfig(**cfg))
    hf_sd = to_transformers_state_dict(sd)
    missing, unexpected = model.load_state_dict(hf_sd, strict=False)
    missing = [k for k in missing if "position_ids" not in k]
    unexpected = [k for k in unexpected if not k.startswith("bert.pooler.")]      # HF's tagger has no pooler
    if missing or unexpected:
        raise RuntimeError("checkpoint does not map onto the transformers model: missing %s, unexpected %s" % (missing[:4], unexpected[:4]))
    return model, (lambda m, batch: m(**batch)["logits"])


def evaluate(model_ckpt, config_file, vocab, test_file, label_list=None, batch_size=32, device=None, impl="native",
             return_logits=False):
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    state = torch.load(model_ckpt, map_location="cpu", weights_only=False)
    sd = state["model"] if "model" in state else state
    sents = conll.read_conll(test_file)
    labels = label_list or getattr(state.get("args", None), "label_list", None) or conll.label_list_from(sents)
    model, forward = _build_model(impl, config_file, len(labels), sd)
    model.to(device).eval()
    all_logits = []
    tok = conll.build_tokenizer(vocab)
    feats = conll.tokenize_and_align(tok, sents, {l: i for i, l in enumerate(labels)})
    coll = YD_DataCollatorForTokenClassification(tok)
    y_true, y_pred = [], []
    with torch.no_grad():
        for s in range(0, len(feats), batch_size):
            b = coll(feats[s:s + batch_size])
            lab = b.pop("labels")
            logits = forward(model, {k: v.to(device) for k, v in b.items()})
            if return_logits:
                all_logits.append(logits.float().cpu())
            pred = logits.argmax(-1).cpu()
            for p_row, l_row in zip(pred.tolist(), lab.tolist()):
                keep = [(labels[p], labels[l]) for p, l in zip(p_row, l_row) if l != -100]
                if keep:
                    y_pred.append([k[0] for k in keep])
                    y_true.append([k[1] for k in keep])
    out = {"accuracy_score": seqeval_lite.accuracy_score(y_true, y_pred), "precision": seqeval_lite.precision_score(y_true, y_pred),
           "recall": seqeval_lite.recall_score(y_true, y_pred), "f1": seqeval_lite.f1_score(y_true, y_pred)}
    print(out)
    if return_logits:
        out["logits"] = all_logits
    return out


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--model_ckpt", required=True)
    ap.add_argument("--config_file", required=True)
    ap.add_argument("--dict", required=True)
    ap.add_argument("--test_file", required=True)
    ap.add_argument("--batch_size", type=int, default=32)
    ap.add_argument("--impl", default="native", choices=["native", "transformers"],
                    help="transformers: load the checkpoint into Hugging Face's BertForTokenClassification instead")
    a = ap.parse_args(argv)
    evaluate(a.model_ckpt, a.config_file, a.dict, a.test_file, batch_size=a.batch_size, impl=a.impl)


if __name__ == "__main__":
    main()
