from .data_collator import YD_DataCollatorForTokenClassification, YD_DataCollatorForELClassification

__all__ = ["YD_DataCollatorForTokenClassification", "YD_DataCollatorForELClassification"]
