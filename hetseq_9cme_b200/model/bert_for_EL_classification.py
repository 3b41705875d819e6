"""BERT for joint mention detection + entity linking (hetseq/model/bert_for_EL_classification.py:21-113).

Encoder -> dropout -> (a) token classifier with masked cross-entropy, (b) ``tanh(Linear(H, dim_entity_emb))``
trained with a cosine-embedding loss against frozen pretrained entity vectors at positions whose
``entity_labels > 0``.  A NaN entity loss (no linked position in the batch) is dropped.  The reference prints
both losses every step; here they are kept on ``self.last_losses`` instead."""
from __future__ import annotations

import torch
import torch.nn as nn
import torch.nn.functional as F

from ..bert_modeling import BertModel, BertPreTrainedModel, _HeadModel


class BertForELClassification(_HeadModel):
    def __init__(self, config, args):
        super().__init__(config)
        self.args = args
        self.num_labels = args.num_labels
        self.bert = BertModel(config)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        self.classifier = nn.Linear(config.hidden_size, self.num_labels)
        self.num_entity_labels = args.num_entity_labels
        self.dim_entity_emb = args.dim_entity_emb
        self.entity_classifier = nn.Linear(config.hidden_size, self.dim_entity_emb)
        self.apply(self.init_bert_weights)
        self.entity_emb = nn.Embedding.from_pretrained(args.EntityEmbedding, freeze=True)
        assert tuple(self.entity_emb.weight.shape) == (self.num_entity_labels, self.dim_entity_emb)
        self.activate = torch.tanh
        self.last_losses = None

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, labels=None, entity_labels=None,
                checkpoint_activations=False):
        seq, _ = self._run_bert(input_ids, token_type_ids, attention_mask, checkpoint_activations, want_pooled=False)
        seq = self.dropout(seq)
        logits = self.classifier(seq)
        entity_logits = self.activate(self.entity_classifier(seq))
        if labels is None:
            return logits, entity_logits
        if attention_mask is None:
            raise ValueError("mask has to not None ")
        active = attention_mask.view(-1) == 1
        lab = torch.where(active, labels.view(-1), torch.full_like(labels.view(-1), -100))
        ner_loss = F.cross_entropy(logits.view(-1, self.num_labels).float(), lab, ignore_index=-100)
        sel = entity_labels.view(-1) > 0
        pred = entity_logits.view(-1, self.dim_entity_emb)[sel]
        gold = self.entity_emb.weight[entity_labels.view(-1)[sel]]
        if pred.numel() == 0:
            entity_loss = torch.full((), float("nan"), device=ner_loss.device)
        else:
            entity_loss = F.cosine_embedding_loss(pred.float(), gold.float(), torch.ones(pred.size(0), device=pred.device))
        self.last_losses = (ner_loss.detach(), entity_loss.detach())
        loss = ner_loss if torch.isnan(entity_loss) else ner_loss + entity_loss
        return loss
