"""Dynamic-padding collators for token classification / entity linking
(API of hetseq/data_collator/data_collator.py:9-310).

Every feature of a batch is padded on the tokenizer's padding side to the longest sequence in the batch:
``input_ids`` with 0, ``labels`` (and ``entity_labels``) with -100, ``token_type_ids`` and ``attention_mask`` with 0;
the result is a dict of int64 tensors.  Implemented with one pre-filled tensor per column instead of Python
list concatenation per sample; ``pad_to_multiple_of`` is honoured (useful: the kernels like multiples of 8).
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Any, Dict, List, Optional, Sequence

import torch


def _as_list(x):
    if torch.is_tensor(x):
        return x.tolist()
    return list(x)


@dataclass
class YD_DataCollatorForTokenClassification:
    tokenizer: Any = None
    padding: Any = True
    max_length: Optional[int] = None
    pad_to_multiple_of: Optional[int] = None
    label_pad_token_id: int = -100

    _NER_COLUMNS = ["input_ids", "labels", "token_type_ids", "attention_mask"]
    INPUT_IDS_PAD = 0
    LABELS_PAD = -100
    TOKEN_TYPE_ID_PAD = 0
    ATTENTION_MASK_PAD = 0

    def _pads(self) -> Dict[str, int]:
        return {"input_ids": self.INPUT_IDS_PAD, "labels": self.LABELS_PAD, "token_type_ids": self.TOKEN_TYPE_ID_PAD,
                "attention_mask": self.ATTENTION_MASK_PAD}

    def __call__(self, features: Sequence[Dict[str, Any]]) -> Dict[str, torch.Tensor]:
        label_name = "label" if "label" in features[0] else "labels"
        width = max(len(f[label_name]) for f in features)
        if self.padding == "max_length" and self.max_length:
            width = max(width, self.max_length)
        if self.pad_to_multiple_of:
            width = (width + self.pad_to_multiple_of - 1) // self.pad_to_multiple_of * self.pad_to_multiple_of
        right = getattr(self.tokenizer, "padding_side", "right") == "right"
        out = {}
        for col, pad in self._pads().items():
            src = label_name if col == "labels" else col
            t = torch.full((len(features), width), pad, dtype=torch.long)
            for r, f in enumerate(features):
                if src not in f:
                    if col == "attention_mask":
                        n = len(f[label_name])
                        vals = [1] * n
                    elif col == "token_type_ids":
                        vals = [0] * len(f[label_name])
                    else:
                        raise KeyError(src)
                else:
                    vals = _as_list(f[src])
                n = len(vals)
                if right:
                    t[r, :n] = torch.as_tensor(vals, dtype=torch.long)
                else:
                    t[r, width - n:] = torch.as_tensor(vals, dtype=torch.long)
            out[col] = t
        return out


@dataclass
class YD_DataCollatorForELClassification(YD_DataCollatorForTokenClassification):
    _NER_COLUMNS = ["input_ids", "labels", "token_type_ids", "attention_mask", "entity_labels"]
    ENTITY_LABELS_PAD = -100

    def _pads(self):
        p = super()._pads()
        p["entity_labels"] = self.ENTITY_LABELS_PAD
        return p
