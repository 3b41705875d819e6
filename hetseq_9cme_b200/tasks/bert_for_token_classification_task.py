"""BERT token-classification (NER) fine-tuning task
(API of hetseq/tasks/bert_for_token_classification_task.py:14-228).

Differences from the reference: CoNLL files are parsed directly (``--extension_file`` with an HF loading script is
honoured when ``datasets`` can run it), all three splits can be loaded (the reference only ever builds 'train'),
and nothing un-picklable is stored on ``args`` that would end up in checkpoints (the checkpoint writer strips the
tokenizer / datasets / collator)."""
from __future__ import annotations

import os

import torch

from ..data import BertNerDataset
from ..data import conll
from ..data_collator import YD_DataCollatorForTokenClassification
from .tasks import Task

_NER_COLUMNS = ["input_ids", "labels", "token_type_ids", "attention_mask"]


class BertForTokenClassificationTask(Task):
    collator_cls = YD_DataCollatorForTokenClassification
    dataset_cls = BertNerDataset

    def __init__(self, args):
        super().__init__(args)
        self._NER_COLUMNS = _NER_COLUMNS

    @classmethod
    def _read_splits(cls, args):
        files = {}
        for split, attr in (("train", "train_file"), ("validation", "validation_file"), ("test", "test_file")):
            f = getattr(args, attr, None)
            if f:
                files[split] = f
        if not files:
            raise ValueError('dataset must contain "train"/"validation"/"test"')
        ext = getattr(args, "extension_file", None)
        if ext and os.path.exists(ext):
            try:
                import datasets  # type: ignore
                ds = datasets.load_dataset(ext, data_files=files)
                first = next(iter(ds.values()))
                names = first.features["ner_tags"].feature.names
                return {k: [(r["tokens"], [names[t] for t in r["ner_tags"]]) for r in v] for k, v in ds.items()}, list(names)
            except Exception as e:  # script needs network / unsupported datasets version -> parse directly
                print("| could not run --extension_file ({}); reading CoNLL files directly".format(type(e).__name__))
        splits = {k: conll.read_conll(v) for k, v in files.items()}
        ref = splits.get("train") or next(iter(splits.values()))
        return splits, conll.label_list_from(ref)

    @classmethod
    def setup_task(cls, args, **kwargs):
        tokenizer = conll.build_tokenizer(args.dict, getattr(args, "max_pred_length", 512))
        collator = cls.collator_cls(tokenizer, max_length=getattr(args, "max_pred_length", 512), padding=True,
                                    pad_to_multiple_of=(getattr(args, "pad_to_multiple_of", 0) or None))
        splits, label_list = cls._read_splits(args)
        label_to_id = {l: i for i, l in enumerate(label_list)}
        tokenized = {k: conll.tokenize_and_align(tokenizer, v, label_to_id, getattr(args, "max_pred_length", 512))
                     for k, v in splits.items()}
        args.tokenized_datasets = tokenized
        args.num_labels = len(label_list)
        args.label_list = label_list
        args.tokenizer = tokenizer
        args.data_collator = collator
        return cls(args)

    def build_model(self, args):
        from ..bert_modeling import BertConfig, BertForTokenClassification
        config = BertConfig.from_json_file(args.config_file)
        model = BertForTokenClassification(config, args.num_labels)
        self._warm_start(model, args)
        return model

    @staticmethod
    def _warm_start(model, args):
        strict = bool(getattr(args, "load_state_dict_strict", False))
        if getattr(args, "hetseq_state_dict", None):
            sd = torch.load(args.hetseq_state_dict, map_location="cpu", weights_only=False)["model"]
            res = model.load_state_dict(sd, strict=strict)
            print("| warm start from {} (missing {}, unexpected {})".format(args.hetseq_state_dict, len(res.missing_keys), len(res.unexpected_keys)))
        elif getattr(args, "transformers_state_dict", None):
            sd = torch.load(args.transformers_state_dict, map_location="cpu", weights_only=False)
            sd = {_hf_key(k): v for k, v in sd.items()}
            res = model.load_state_dict(sd, strict=strict)
            print("| warm start from {} (missing {}, unexpected {})".format(args.transformers_state_dict, len(res.missing_keys), len(res.unexpected_keys)))

    def load_dataset(self, split, **kwargs):
        key = {"valid": "validation"}.get(split, split)
        data = self.args.tokenized_datasets
        if key not in data:
            # every shipped script passes --valid-subset test although validation never ran; be lenient
            key = next(iter(data))
        self.datasets[split] = self.dataset_cls(data[key], self.args)
        print("| loaded {} sentences for split {}".format(len(self.datasets[split]), split), flush=True)

    def _forward(self, sample, model):
        return model(**sample)

    def _sample_size(self, sample) -> int:
        if sample is None or len(sample) == 0:
            return 0
        return len(sample["labels"])           # batch size (ref: :212-215)

    def _num_tokens(self, sample) -> int:
        try:
            return int(sample["attention_mask"].numel())
        except Exception:
            return 0


def _hf_key(k: str) -> str:
    """HF transformers BERT key -> this model's key."""
    k = k.replace("intermediate.dense.", "intermediate.dense_act.").replace("pooler.dense.", "pooler.dense_act.")
    k = k.replace("cls.predictions.transform.dense.", "cls.predictions.transform.dense_act.")
    k = k.replace("LayerNorm.gamma", "LayerNorm.weight").replace("LayerNorm.beta", "LayerNorm.bias")
    return k
