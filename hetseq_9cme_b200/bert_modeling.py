"""BERT model family.

Same module tree, parameter names and forward signatures as the reference model file
(hetseq/bert_modeling.py:132-1329) so checkpoints are interchangeable, with two execution paths:

* **native** (CUDA + parameters bound to a ``ParamArena``): embeddings, every encoder layer and the
  pre-training / token-classification heads run as hand-scheduled sm_100a kernel sequences
  (``models/bert_native.py``).  Selected automatically; there is no silent fallback on a GPU unless the
  configuration is outside what the kernels cover, in which case a one-line notice is printed.
* **torch** (CPU, debugging, unsupported shapes): plain PyTorch ops written directly from the math.

Models return the loss when labels are given (contract documented in the reference's
docs/source/extending.rst:52), logits otherwise.
"""
from __future__ import annotations

import copy
import json
import math
from typing import List, Optional

import torch
import torch.nn as nn
import torch.nn.functional as F


# -------------------------------------------------------------------------------------------------
# config
# -------------------------------------------------------------------------------------------------
class BertConfig:
    """Hyper-parameters of a BERT encoder; reads the same ``bert_config.json`` keys as the reference
    (hetseq/bert_modeling.py:180-266)."""

    _DEFAULTS = dict(hidden_size=768, num_hidden_layers=12, num_attention_heads=12, intermediate_size=3072,
                     hidden_act="gelu", hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1,
                     max_position_embeddings=512, type_vocab_size=2, initializer_range=0.02)

    def __init__(self, vocab_size_or_config_json_file=30522, **kwargs):
        for k, v in self._DEFAULTS.items():
            setattr(self, k, kwargs.pop(k, v))
        if kwargs:
            for k, v in kwargs.items():
                setattr(self, k, v)
        if isinstance(vocab_size_or_config_json_file, str):
            with open(vocab_size_or_config_json_file, "r", encoding="utf-8") as f:
                self.__dict__.update(json.load(f))
        elif isinstance(vocab_size_or_config_json_file, int):
            self.vocab_size = vocab_size_or_config_json_file
        else:
            raise ValueError("First argument must be either a vocabulary size (int) or the path to a config file (str)")

    @classmethod
    def from_dict(cls, json_object):
        cfg = cls(vocab_size_or_config_json_file=-1)
        cfg.__dict__.update(json_object)
        return cfg

    @classmethod
    def from_json_file(cls, json_file):
        with open(json_file, "r", encoding="utf-8") as f:
            return cls.from_dict(json.load(f))

    def to_dict(self):
        return copy.deepcopy(self.__dict__)

    def to_json_string(self):
        return json.dumps(self.to_dict(), indent=2, sort_keys=True) + "\n"

    def __repr__(self):
        return self.to_json_string()


# -------------------------------------------------------------------------------------------------
# activations and small layers
# -------------------------------------------------------------------------------------------------
def gelu(x):
    return x * 0.5 * (1.0 + torch.erf(x / math.sqrt(2.0)))


def swish(x):
    return x * torch.sigmoid(x)


ACT2FN = {"gelu": gelu, "relu": F.relu, "swish": swish, "tanh": torch.tanh, "noact": lambda x: x}


class LinearActivation(nn.Module):
    """``act(x W^T + b)`` with the activation fused into the GEMM epilogue on the native path
    (ref: hetseq/bert_modeling.py:132-177)."""

    def __init__(self, in_features, out_features, act="gelu", bias=True):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.act_name = act if isinstance(act, str) else getattr(act, "__name__", "custom")
        self.act_fn = ACT2FN[act] if isinstance(act, str) else act
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.empty(out_features)) if bias else None
        self.reset_parameters()

    def reset_parameters(self):
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        if self.bias is not None:
            bound = 1.0 / math.sqrt(self.in_features)
            nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x):
        return self.act_fn(F.linear(x, self.weight, self.bias))

    def extra_repr(self):
        return "in_features={}, out_features={}, act={}".format(self.in_features, self.out_features, self.act_name)


class BertLayerNorm(nn.Module):
    """TF-style LayerNorm: biased variance, epsilon inside the square root (ref: :276-289)."""

    def __init__(self, hidden_size, eps=1e-12):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(hidden_size))
        self.bias = nn.Parameter(torch.zeros(hidden_size))
        self.variance_epsilon = eps

    def forward(self, x):
        xf = x.float()
        mean = xf.mean(-1, keepdim=True)
        var = (xf - mean).pow(2).mean(-1, keepdim=True)
        xn = ((xf - mean) * torch.rsqrt(var + self.variance_epsilon)).to(x.dtype)
        return self.weight * xn + self.bias


class BertEmbeddings(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.word_embeddings = nn.Embedding(config.vocab_size, config.hidden_size)
        self.position_embeddings = nn.Embedding(config.max_position_embeddings, config.hidden_size)
        self.token_type_embeddings = nn.Embedding(config.type_vocab_size, config.hidden_size)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=1e-12)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)

    def forward(self, input_ids, token_type_ids=None):
        L = input_ids.size(1)
        pos = torch.arange(L, dtype=torch.long, device=input_ids.device).unsqueeze(0).expand_as(input_ids)
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        e = self.word_embeddings(input_ids) + self.position_embeddings(pos) + self.token_type_embeddings(token_type_ids)
        return self.dropout(self.LayerNorm(e))


class BertSelfAttention(nn.Module):
    def __init__(self, config):
        super().__init__()
        if config.hidden_size % config.num_attention_heads != 0:
            raise ValueError("hidden size (%d) is not a multiple of the number of attention heads (%d)"
                             % (config.hidden_size, config.num_attention_heads))
        self.num_attention_heads = config.num_attention_heads
        self.attention_head_size = config.hidden_size // config.num_attention_heads
        self.all_head_size = self.num_attention_heads * self.attention_head_size
        self.query = nn.Linear(config.hidden_size, self.all_head_size)
        self.key = nn.Linear(config.hidden_size, self.all_head_size)
        self.value = nn.Linear(config.hidden_size, self.all_head_size)
        self.dropout = nn.Dropout(config.attention_probs_dropout_prob)

    def _heads(self, x):
        B, L, _ = x.shape
        return x.view(B, L, self.num_attention_heads, self.attention_head_size).transpose(1, 2)

    def forward(self, hidden_states, attention_mask):
        q, k, v = self._heads(self.query(hidden_states)), self._heads(self.key(hidden_states)), self._heads(self.value(hidden_states))
        scores = torch.matmul(q, k.transpose(-1, -2)) / math.sqrt(self.attention_head_size)
        if attention_mask is not None:
            scores = scores + attention_mask
        probs = self.dropout(torch.softmax(scores, dim=-1))
        ctx = torch.matmul(probs, v).transpose(1, 2).contiguous()
        return ctx.view(ctx.size(0), ctx.size(1), self.all_head_size)


class BertSelfOutput(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.hidden_size, config.hidden_size)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=1e-12)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)

    def forward(self, hidden_states, input_tensor):
        return self.LayerNorm(self.dropout(self.dense(hidden_states)) + input_tensor)


class BertAttention(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.self = BertSelfAttention(config)
        self.output = BertSelfOutput(config)

    def forward(self, input_tensor, attention_mask):
        return self.output(self.self(input_tensor, attention_mask), input_tensor)


class BertIntermediate(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense_act = LinearActivation(config.hidden_size, config.intermediate_size, act=config.hidden_act)

    def forward(self, hidden_states):
        return self.dense_act(hidden_states)


class BertOutput(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense = nn.Linear(config.intermediate_size, config.hidden_size)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=1e-12)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)

    def forward(self, hidden_states, input_tensor):
        return self.LayerNorm(self.dropout(self.dense(hidden_states)) + input_tensor)


class BertLayer(nn.Module):
    """Post-LN transformer block."""

    def __init__(self, config):
        super().__init__()
        self.attention = BertAttention(config)
        self.intermediate = BertIntermediate(config)
        self.output = BertOutput(config)

    def forward(self, hidden_states, attention_mask):
        a = self.attention(hidden_states, attention_mask)
        return self.output(self.intermediate(a), a)


class BertEncoder(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.layer = nn.ModuleList([BertLayer(config) for _ in range(config.num_hidden_layers)])

    def forward(self, hidden_states, attention_mask, output_all_encoded_layers=True, checkpoint_activations=False):
        outs = []
        if checkpoint_activations and self.training:
            # sqrt(N)-sized chunks recomputed in backward (ref: :459-476)
            from torch.utils.checkpoint import checkpoint
            n = len(self.layer)
            chunk = max(1, int(math.ceil(math.sqrt(n))))

            def run(start, end):
                def fn(h, m):
                    for lyr in self.layer[start:end]:
                        h = lyr(h, m)
                    return h
                return fn

            i = 0
            while i < n:
                hidden_states = checkpoint(run(i, min(n, i + chunk)), hidden_states, attention_mask, use_reentrant=False)
                i += chunk
        else:
            for lyr in self.layer:
                hidden_states = lyr(hidden_states, attention_mask)
                if output_all_encoded_layers:
                    outs.append(hidden_states)
        if not output_all_encoded_layers or checkpoint_activations:
            outs.append(hidden_states)
        return outs


class BertPooler(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense_act = LinearActivation(config.hidden_size, config.hidden_size, act="tanh")

    def forward(self, hidden_states):
        return self.dense_act(hidden_states[:, 0])


class BertPredictionHeadTransform(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.dense_act = LinearActivation(config.hidden_size, config.hidden_size, act=config.hidden_act)
        self.LayerNorm = BertLayerNorm(config.hidden_size, eps=1e-12)

    def forward(self, hidden_states):
        return self.LayerNorm(self.dense_act(hidden_states))


class BertLMPredictionHead(nn.Module):
    """Decoder weight is tied to the word embeddings; only the output bias is its own parameter."""

    def __init__(self, config, bert_model_embedding_weights):
        super().__init__()
        self.transform = BertPredictionHeadTransform(config)
        self.decoder = nn.Linear(bert_model_embedding_weights.size(1), bert_model_embedding_weights.size(0), bias=False)
        self.decoder.weight = bert_model_embedding_weights
        self.bias = nn.Parameter(torch.zeros(bert_model_embedding_weights.size(0)))

    def forward(self, hidden_states):
        return self.decoder(self.transform(hidden_states)) + self.bias


class BertOnlyMLMHead(nn.Module):
    def __init__(self, config, bert_model_embedding_weights):
        super().__init__()
        self.predictions = BertLMPredictionHead(config, bert_model_embedding_weights)

    def forward(self, sequence_output):
        return self.predictions(sequence_output)


class BertOnlyNSPHead(nn.Module):
    def __init__(self, config):
        super().__init__()
        self.seq_relationship = nn.Linear(config.hidden_size, 2)

    def forward(self, pooled_output):
        return self.seq_relationship(pooled_output)


class BertPreTrainingHeads(nn.Module):
    def __init__(self, config, bert_model_embedding_weights):
        super().__init__()
        self.predictions = BertLMPredictionHead(config, bert_model_embedding_weights)
        self.seq_relationship = nn.Linear(config.hidden_size, 2)

    def forward(self, sequence_output, pooled_output):
        return self.predictions(sequence_output), self.seq_relationship(pooled_output)


# -------------------------------------------------------------------------------------------------
# base class
# -------------------------------------------------------------------------------------------------
class BertPreTrainedModel(nn.Module):
    """Weight initialisation + native-engine plumbing shared by every head model."""

    def __init__(self, config, *inputs, **kwargs):
        super().__init__()
        if not isinstance(config, BertConfig):
            raise ValueError("Parameter config should be an instance of BertConfig")
        self.config = config
        self._arena = None
        self._grad_sink = None
        self._native_notice = False

    def init_bert_weights(self, module):
        """N(0, initializer_range) for Linear / Embedding weights, LayerNorm = (1, 0), biases 0 (ref: :599-610)."""
        if isinstance(module, (nn.Linear, nn.Embedding, LinearActivation)):
            module.weight.data.normal_(mean=0.0, std=self.config.initializer_range)
        elif isinstance(module, BertLayerNorm):
            module.bias.data.zero_()
            module.weight.data.fill_(1.0)
        if isinstance(module, (nn.Linear, LinearActivation)) and module.bias is not None:
            module.bias.data.zero_()

    # ---- native engine ----------------------------------------------------------------------------
    def arena_groups(self):
        """Parameters that must be adjacent in the arena: per layer (q,k,v) weights and (q,k,v) biases."""
        groups = []
        bert = self if isinstance(self, BertModel) else getattr(self, "bert", None)
        if bert is not None:
            for lyr in bert.encoder.layer:
                s = lyr.attention.self
                groups.append([s.query.weight, s.key.weight, s.value.weight])
                groups.append([s.query.bias, s.key.bias, s.value.bias])
        return groups

    def bind_arena(self, arena, grad_sink=None):
        """Called by the Controller once parameters live in a ParamArena: builds the per-layer packed views."""
        from .models import bert_native
        self._arena = arena
        self._grad_sink = grad_sink
        bert = self if isinstance(self, BertModel) else getattr(self, "bert", None)
        if bert is not None and arena is not None and arena.device.type == "cuda":
            bert._native_packs = [bert_native.LayerPack(lyr, arena) for lyr in bert.encoder.layer]

    def _native_ctx(self, input_ids, force_torch: bool = False):
        """A NativeCtx when the kernels should run this call, else None."""
        if force_torch or self._arena is None or not input_ids.is_cuda or getattr(self, "force_torch_path", False):
            return None
        from .models import bert_native
        why = bert_native.native_supported(self.config, input_ids.size(1))
        if why is not None:
            # an arena-bound model on a GPU is expected to run the sm_100a kernels: dropping a whole model to eager
            # PyTorch must be the user's explicit choice (--no-native / force_torch_path), never a quiet side effect
            if not getattr(self, "allow_torch_fallback", False):
                raise RuntimeError("the native kernels do not cover this model ({}); pass --no-native to run it on the "
                                   "plain PyTorch path".format(why))
            if not self._native_notice:
                print("| native kernels disabled for this model: {}".format(why), flush=True)
                self._native_notice = True
            return None
        return bert_native.NativeCtx(input_ids.device, self.training, self._grad_sink,
                                     checkpoint_activations=getattr(self, "checkpoint_activations", False) and self.training)

    @classmethod
    def from_pretrained(cls, path, *inputs, **kwargs):
        """Load ``bert_config.json`` + ``pytorch_model.bin`` from a local directory.  (The reference also
        downloads named archives from S3, hetseq/bert_modeling.py:612-752; there is no network here.)"""
        import os
        config = BertConfig.from_json_file(os.path.join(path, "bert_config.json"))
        model = cls(config, *inputs, **kwargs)
        wpath = os.path.join(path, "pytorch_model.bin")
        if os.path.exists(wpath):
            sd = torch.load(wpath, map_location="cpu", weights_only=False)
            sd = {k.replace("gamma", "weight").replace("beta", "bias"): v for k, v in sd.items()}
            target = model if any(k.startswith("bert.") for k in sd) or not hasattr(model, "bert") else model.bert
            missing, unexpected = target.load_state_dict(sd, strict=False)
            if missing:
                print("| from_pretrained: missing keys: {}".format(len(missing)))
        return model


def _extended_mask(attention_mask, dtype):
    """(1 - mask) * -10000 broadcast over heads and queries (ref: :817-825)."""
    ext = attention_mask[:, None, None, :].to(dtype)
    return (1.0 - ext) * -10000.0


class BertModel(BertPreTrainedModel):
    def __init__(self, config):
        super().__init__(config)
        self.embeddings = BertEmbeddings(config)
        self.encoder = BertEncoder(config)
        self.pooler = BertPooler(config)
        self.apply(self.init_bert_weights)
        self._native_packs = None

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, output_all_encoded_layers=True,
                checkpoint_activations=False):
        if attention_mask is None:
            attention_mask = torch.ones_like(input_ids)
        if token_type_ids is None:
            token_type_ids = torch.zeros_like(input_ids)
        emb = self.embeddings(input_ids, token_type_ids)
        ext = _extended_mask(attention_mask, emb.dtype)
        layers = self.encoder(emb, ext, output_all_encoded_layers=output_all_encoded_layers,
                              checkpoint_activations=checkpoint_activations)
        pooled = self.pooler(layers[-1])
        if not output_all_encoded_layers:
            layers = layers[-1]
        return layers, pooled

    def native_forward(self, nctx, input_ids, token_type_ids, attention_mask, want_pooled=True):
        """[B*L, H] final hidden states (+ pooled output) through the sm_100a kernels."""
        from .models import bert_native
        B, L = input_ids.shape
        seq2d = bert_native.encoder_forward(self, nctx, input_ids, token_type_ids, attention_mask)
        pooled = bert_native.PoolerFn.apply(seq2d, nctx, self.pooler, B, L) if want_pooled else None
        return seq2d, pooled


class _HeadModel(BertPreTrainedModel):
    """Shared helpers for models of the form ``self.bert`` + head."""

    def _run_bert(self, input_ids, token_type_ids, attention_mask, checkpoint_activations=False, want_pooled=True):
        nctx = self._native_ctx(input_ids)
        B, L = input_ids.shape
        if nctx is not None:
            seq2d, pooled = self.bert.native_forward(nctx, input_ids, token_type_ids, attention_mask, want_pooled)
            return seq2d.view(B, L, -1), pooled
        seq, pooled = self.bert(input_ids, token_type_ids, attention_mask, output_all_encoded_layers=False,
                                checkpoint_activations=checkpoint_activations or getattr(self, "checkpoint_activations", False))
        return seq, pooled

    def bind_arena(self, arena, grad_sink=None):
        super().bind_arena(arena, grad_sink)
        self.bert._arena = arena
        self.bert._grad_sink = grad_sink


class BertForPreTraining(_HeadModel):
    """Masked-LM + next-sentence heads; ``forward`` returns the summed loss when both label tensors are given."""

    def __init__(self, config):
        super().__init__(config)
        self.bert = BertModel(config)
        self.cls = BertPreTrainingHeads(config, self.bert.embeddings.word_embeddings.weight)
        self.apply(self.init_bert_weights)
        self.max_predictions_per_seq = 0     # 0: no bound known -> capacity = sequence length
        self.last_mlm_count = None

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, masked_lm_labels=None,
                next_sentence_label=None, checkpoint_activations=False):
        nctx = self._native_ctx(input_ids)
        if nctx is not None and masked_lm_labels is not None and next_sentence_label is not None:
            from .models import bert_native
            B, L = input_ids.shape
            seq2d = bert_native.encoder_forward(self.bert, nctx, input_ids, token_type_ids, attention_mask)
            per_seq = self.max_predictions_per_seq if 0 < self.max_predictions_per_seq <= L else L
            cap = (B * per_seq + 3) // 4 * 4
            loss, cnt = bert_native.PretrainLossFn.apply(seq2d, nctx, self, masked_lm_labels, next_sentence_label, B, L, cap)
            self.last_mlm_count = cnt     # device int32 [2]: (used, found); found > used means capacity overflow
            return loss
        seq, pooled = self._run_bert(input_ids, token_type_ids, attention_mask, checkpoint_activations)
        scores, nsp = self.cls(seq, pooled)
        if masked_lm_labels is not None and next_sentence_label is not None:
            mlm = F.cross_entropy(scores.view(-1, self.config.vocab_size).float(), masked_lm_labels.view(-1), ignore_index=-1)
            nsl = F.cross_entropy(nsp.view(-1, 2).float(), next_sentence_label.view(-1), ignore_index=-1)
            return mlm + nsl
        return scores, nsp


class BertForMaskedLM(_HeadModel):
    def __init__(self, config):
        super().__init__(config)
        self.bert = BertModel(config)
        self.cls = BertOnlyMLMHead(config, self.bert.embeddings.word_embeddings.weight)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, masked_lm_labels=None, checkpoint_activations=False):
        seq, _ = self._run_bert(input_ids, token_type_ids, attention_mask, checkpoint_activations, want_pooled=False)
        scores = self.cls(seq)
        if masked_lm_labels is not None:
            return F.cross_entropy(scores.view(-1, self.config.vocab_size).float(), masked_lm_labels.view(-1), ignore_index=-1)
        return scores


class BertForNextSentencePrediction(_HeadModel):
    def __init__(self, config):
        super().__init__(config)
        self.bert = BertModel(config)
        self.cls = BertOnlyNSPHead(config)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, next_sentence_label=None, checkpoint_activations=False):
        _, pooled = self._run_bert(input_ids, token_type_ids, attention_mask, checkpoint_activations)
        score = self.cls(pooled)
        if next_sentence_label is not None:
            return F.cross_entropy(score.view(-1, 2).float(), next_sentence_label.view(-1), ignore_index=-1)
        return score


class BertForSequenceClassification(_HeadModel):
    def __init__(self, config, num_labels):
        super().__init__(config)
        self.num_labels = num_labels
        self.bert = BertModel(config)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        self.classifier = nn.Linear(config.hidden_size, num_labels)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, labels=None, checkpoint_activations=False):
        _, pooled = self._run_bert(input_ids, token_type_ids, attention_mask, checkpoint_activations)
        logits = self.classifier(self.dropout(pooled))
        if labels is not None:
            return F.cross_entropy(logits.view(-1, self.num_labels).float(), labels.view(-1))
        return logits


class BertForMultipleChoice(_HeadModel):
    def __init__(self, config, num_choices):
        super().__init__(config)
        self.num_choices = num_choices
        self.bert = BertModel(config)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        self.classifier = nn.Linear(config.hidden_size, 1)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, labels=None, checkpoint_activations=False):
        flat = lambda t: t.view(-1, t.size(-1)) if t is not None else None
        _, pooled = self._run_bert(flat(input_ids), flat(token_type_ids), flat(attention_mask), checkpoint_activations)
        logits = self.classifier(self.dropout(pooled)).view(-1, self.num_choices)
        if labels is not None:
            return F.cross_entropy(logits.float(), labels)
        return logits


class BertForTokenClassification(_HeadModel):
    """Per-token classifier (NER); loss over positions with ``attention_mask == 1`` and label != -100."""

    def __init__(self, config, num_labels):
        super().__init__(config)
        self.num_labels = num_labels
        self.bert = BertModel(config)
        self.dropout = nn.Dropout(config.hidden_dropout_prob)
        self.classifier = nn.Linear(config.hidden_size, num_labels)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, labels=None, checkpoint_activations=False):
        nctx = self._native_ctx(input_ids)
        B, L = input_ids.shape
        if nctx is not None and self.num_labels <= 32:
            from .models import bert_native
            seq2d = bert_native.encoder_forward(self.bert, nctx, input_ids, token_type_ids, attention_mask)
            if labels is None:
                from .ops import kernels as K
                return K.head_fwd(seq2d.detach(), self.classifier.weight, self.classifier.bias).view(B, L, -1)
            loss, logits = bert_native.TokenClsLossFn.apply(seq2d, nctx, self, labels, attention_mask, 7 << 20)
            self.last_logits = logits.view(B, L, -1)
            return loss
        seq, _ = self._run_bert(input_ids, token_type_ids, attention_mask, checkpoint_activations, want_pooled=False)
        logits = self.classifier(self.dropout(seq))
        if labels is not None:
            lab = labels.view(-1)
            if attention_mask is not None:
                lab = torch.where(attention_mask.view(-1) == 1, lab, torch.full_like(lab, -100))
            return F.cross_entropy(logits.view(-1, self.num_labels).float(), lab, ignore_index=-100)
        return logits


class BertForQuestionAnswering(_HeadModel):
    def __init__(self, config):
        super().__init__(config)
        self.bert = BertModel(config)
        self.qa_outputs = nn.Linear(config.hidden_size, 2)
        self.apply(self.init_bert_weights)

    def forward(self, input_ids, token_type_ids=None, attention_mask=None, start_positions=None, end_positions=None,
                checkpoint_activations=False):
        seq, _ = self._run_bert(input_ids, token_type_ids, attention_mask, checkpoint_activations, want_pooled=False)
        start_logits, end_logits = self.qa_outputs(seq).unbind(-1)
        if start_positions is not None and end_positions is not None:
            ignored = start_logits.size(1)
            sp = start_positions.view(-1).clamp(0, ignored)
            ep = end_positions.view(-1).clamp(0, ignored)
            return (F.cross_entropy(start_logits.float(), sp, ignore_index=ignored)
                    + F.cross_entropy(end_logits.float(), ep, ignore_index=ignored)) / 2
        return start_logits, end_logits
