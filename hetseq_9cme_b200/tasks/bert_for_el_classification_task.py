"""BERT entity-linking task (API of hetseq/tasks/bert_for_el_classification_task.py:27-304).

Same pipeline as the NER task plus an ``entity_labels`` column.  The reference maps entity names to ids through
the external, non-vendored ``deep_ed_PyTorch`` package and loads entity vectors from ``--ent_vecs_filename``;
here the mapping comes from a plain ``entity_vocab.txt`` (one entity name per line, line number = id, id 0 =
"no entity") next to the entity vectors, and ``deep_ed_PyTorch`` is used only if it happens to be importable.
CoNLL lines may carry the entity name as an extra last column; without it every token is unlinked."""
from __future__ import annotations

import os

import torch

from ..data import BertELDataset, conll
from ..data_collator import YD_DataCollatorForELClassification
from .bert_for_token_classification_task import BertForTokenClassificationTask

_EL_COLUMNS = ["input_ids", "labels", "token_type_ids", "attention_mask", "entity_labels"]


class BertForELClassificationTask(BertForTokenClassificationTask):
    collator_cls = YD_DataCollatorForELClassification
    dataset_cls = BertELDataset

    @classmethod
    def setup_task(cls, args, **kwargs):
        tokenizer = conll.build_tokenizer(args.dict, getattr(args, "max_pred_length", 512))
        collator = cls.collator_cls(tokenizer, max_length=getattr(args, "max_pred_length", 512), padding=True,
                                    pad_to_multiple_of=(getattr(args, "pad_to_multiple_of", 0) or None))
        ent_vecs = torch.load(args.ent_vecs_filename, map_location="cpu", weights_only=False)
        ent_vecs = torch.as_tensor(ent_vecs).float()
        vocab_path = os.path.join(os.path.dirname(args.ent_vecs_filename), "entity_vocab.txt")
        ent2id = {}
        if os.path.exists(vocab_path):
            with open(vocab_path, "r", encoding="utf-8") as f:
                ent2id = {line.strip(): i for i, line in enumerate(f)}
        files = {s: getattr(args, a) for s, a in (("train", "train_file"), ("validation", "validation_file"), ("test", "test_file"))
                 if getattr(args, a, None)}
        splits, ent_cols = {}, {}
        for k, path in files.items():
            sents = conll.read_conll(path, tag_col=3 if _ncols(path) >= 5 else -1)
            splits[k] = sents
            ent_cols[k] = _read_entity_column(path) if _ncols(path) >= 5 else [["O"] * len(t) for t, _ in sents]
        label_list = conll.label_list_from(splits.get("train") or next(iter(splits.values())))
        label_to_id = {l: i for i, l in enumerate(label_list)}
        tokenized = {}
        for k, sents in splits.items():
            feats = conll.tokenize_and_align(tokenizer, sents, label_to_id, getattr(args, "max_pred_length", 512))
            for f, ents in zip(feats, ent_cols[k]):
                word = -1
                el = []
                for lab in f["labels"]:
                    if lab == -100:
                        el.append(-100)
                    else:
                        word += 1
                        el.append(ent2id.get(ents[word], 0) if word < len(ents) else 0)
                f["entity_labels"] = el
            tokenized[k] = feats
        args.tokenized_datasets = tokenized
        args.num_labels = len(label_list)
        args.label_list = label_list
        args.tokenizer = tokenizer
        args.data_collator = collator
        args.EntityEmbedding = ent_vecs
        args.num_entity_labels, args.dim_entity_emb = ent_vecs.shape
        return cls(args)

    def build_model(self, args):
        from ..bert_modeling import BertConfig
        from ..model import BertForELClassification
        config = BertConfig.from_json_file(args.config_file)
        model = BertForELClassification(config, args)
        self._warm_start(model, args)
        return model


def _ncols(path: str) -> int:
    with open(path, "r", encoding="utf-8") as f:
        for line in f:
            if line.strip() and not line.startswith("-DOCSTART-"):
                return len(line.split())
    return 0


def _read_entity_column(path: str):
    out, cur = [], []
    with open(path, "r", encoding="utf-8") as f:
        for line in f:
            if not line.strip() or line.startswith("-DOCSTART-"):
                if cur:
                    out.append(cur)
                    cur = []
                continue
            cur.append(line.split()[-1])
    if cur:
        out.append(cur)
    return out
