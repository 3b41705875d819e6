"""Entity-linking dataset wrapper (module path of hetseq/data/bert_el_dataset.py:7-34).

The class itself lives next to the NER wrapper, which it extends by the ``entity_labels`` feature; this module keeps
the reference's import path ``data.bert_el_dataset.BertELDataset`` working."""
from .bert_ner_dataset import BertELDataset

__all__ = ["BertELDataset"]
