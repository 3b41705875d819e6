from .tasks import Task, LanguageModelingTask, MNISTTask, MNISTNet
from .bert_for_token_classification_task import BertForTokenClassificationTask
from .bert_for_el_classification_task import BertForELClassificationTask

TASK_REGISTRY = {
    "bert": LanguageModelingTask,
    "mnist": MNISTTask,
    "BertForTokenClassification": BertForTokenClassificationTask,
    "BertForELClassification": BertForELClassificationTask,
}

__all__ = ["Task", "LanguageModelingTask", "MNISTTask", "MNISTNet", "BertForTokenClassificationTask",
           "BertForELClassificationTask", "TASK_REGISTRY"]
