"""hetseq_9cme_b200 — a Blackwell-native (sm_100a) data-parallel training framework with the capabilities,
CLI, checkpoint format and task/dataset/model extension contract of HetSeq.

Module names mirror the reference package (``train``, ``options``, ``controller``, ``tasks``, ``optim``,
``lr_scheduler``, ``checkpoint_utils``, ``distributed_utils``, ``meters``, ``progress_bar``, ``utils``, ``data``,
``data_collator``, ``bert_modeling``, ``model``); the sm_100a kernels live in ``csrc/`` behind ``ops``."""

__version__ = "0.1.0"
