"""Token-classification task variant that trains a Hugging Face ``transformers.BertForTokenClassification`` instead
of this framework's own BERT (role of the reference's alternate module hetseq/transformers_tasks.py:395-477:
``model(**sample)['loss']``, ``sample_size = 1``).  The model runs on stock PyTorch ops; gradients still land in the
flat arena, so the fused optimizer step and the peer-to-peer reducer are used unchanged.  Import is lazy: the module
loads without ``transformers``."""
from __future__ import annotations

from .tasks.bert_for_token_classification_task import BertForTokenClassificationTask


class _LossOnly(object):
    pass


class TransformersBertForTokenClassificationTask(BertForTokenClassificationTask):
    def build_model(self, args):
        import json
        import torch
        from transformers import BertConfig as HFConfig, BertForTokenClassification as HFModel
        with open(args.config_file) as f:
            cfg = json.load(f)
        cfg["num_labels"] = args.num_labels
        model = HFModel(HFConfig(**cfg))
        if getattr(args, "transformers_state_dict", None):
            sd = torch.load(args.transformers_state_dict, map_location="cpu", weights_only=False)
            model.load_state_dict(sd, strict=bool(getattr(args, "load_state_dict_strict", False)))
        return model

    def _forward(self, sample, model):
        return model(**sample)["loss"]

    def _sample_size(self, sample) -> int:
        return 0 if sample is None or len(sample) == 0 else 1
