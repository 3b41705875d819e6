from .h5pyDataset import BertH5pyData, ConBertH5pyData
from .mnist_dataset import MNISTDataset
from .bert_ner_dataset import BertNerDataset, BertELDataset
from . import data_utils, iterators

__all__ = ["BertH5pyData", "ConBertH5pyData", "MNISTDataset", "BertNerDataset", "BertELDataset", "data_utils", "iterators"]
