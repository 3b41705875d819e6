from .bert_for_EL_classification import BertForELClassification

__all__ = ["BertForELClassification"]
