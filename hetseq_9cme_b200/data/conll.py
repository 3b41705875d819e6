"""CoNLL-2003 style reader + WordPiece tokenisation/label alignment for the token-classification tasks.

The reference goes through a user-supplied HF ``datasets`` loading script, ``BertTokenizerFast`` and
``dataset.map`` (hetseq/tasks/bert_for_token_classification_task.py:19-131).  Neither network nor the script are
available here, so the files are parsed directly; HF's tokenizer is used when importable, otherwise a small
greedy longest-match WordPiece written for this purpose.  Alignment rule is the reference's: the first sub-token
of a word carries the word's label, continuation sub-tokens and special tokens carry -100.
"""
from __future__ import annotations

import collections
from typing import Dict, List, Optional, Sequence, Tuple


def read_conll(path: str, token_col: int = 0, tag_col: int = -1) -> List[Tuple[List[str], List[str]]]:
    sents, toks, tags = [], [], []
    with open(path, "r", encoding="utf-8") as f:
        for line in f:
            line = line.rstrip("\n")
            if not line.strip() or line.startswith("-DOCSTART-"):
                if toks:
                    sents.append((toks, tags))
                    toks, tags = [], []
                continue
            parts = line.split()
            toks.append(parts[token_col])
            tags.append(parts[tag_col])
    if toks:
        sents.append((toks, tags))
    return sents


def label_list_from(sentences: Sequence[Tuple[List[str], List[str]]]) -> List[str]:
    labs = sorted({t for _, tags in sentences for t in tags})
    if "O" in labs:                       # conventional order: O first
        labs.remove("O")
        labs = ["O"] + labs
    return labs


def load_vocab(path: str) -> "collections.OrderedDict[str, int]":
    vocab = collections.OrderedDict()
    with open(path, "r", encoding="utf-8") as f:
        for i, line in enumerate(f):
            vocab[line.rstrip("\n").strip()] = i
    return vocab


class WordPiece:
    """Minimal uncased BERT tokenizer: lower-case, greedy longest-match-first word pieces."""

    padding_side = "right"

    def __init__(self, vocab_file: str, do_lower_case: bool = True, max_len: int = 512):
        self.vocab = load_vocab(vocab_file)
        self.do_lower_case = do_lower_case
        self.max_len = max_len
        self.unk, self.cls, self.sep, self.pad = (self.vocab.get(t, 0) for t in ("[UNK]", "[CLS]", "[SEP]", "[PAD]"))

    def word_pieces(self, word: str) -> List[int]:
        w = word.lower() if self.do_lower_case else word
        if w in self.vocab:
            return [self.vocab[w]]
        out, start = [], 0
        while start < len(w):
            end, cur = len(w), None
            while start < end:
                piece = w[start:end] if start == 0 else "##" + w[start:end]
                if piece in self.vocab:
                    cur = self.vocab[piece]
                    break
                end -= 1
            if cur is None:
                return [self.unk]
            out.append(cur)
            start = end
        return out

    def encode_words(self, words: Sequence[str], labels: Sequence[int]) -> Dict[str, List[int]]:
        ids, labs = [self.cls], [-100]
        for w, l in zip(words, labels):
            pcs = self.word_pieces(w)
            ids.extend(pcs)
            labs.extend([l] + [-100] * (len(pcs) - 1))
        ids, labs = ids[: self.max_len - 1], labs[: self.max_len - 1]
        ids.append(self.sep)
        labs.append(-100)
        return {"input_ids": ids, "labels": labs, "token_type_ids": [0] * len(ids), "attention_mask": [1] * len(ids)}


def build_tokenizer(vocab_file: str, max_len: int = 512):
    try:
        from transformers import BertTokenizerFast  # type: ignore
        tok = BertTokenizerFast(vocab_file)
        tok.model_max_length = max_len
        return tok
    except Exception:
        return WordPiece(vocab_file, max_len=max_len)


def tokenize_and_align(tokenizer, sentences, label_to_id: Dict[str, int], max_len: int = 512) -> List[Dict[str, List[int]]]:
    feats = []
    if isinstance(tokenizer, WordPiece):
        for words, tags in sentences:
            feats.append(tokenizer.encode_words(words, [label_to_id[t] for t in tags]))
        return feats
    enc = tokenizer([w for w, _ in sentences], padding=False, truncation=True, max_length=max_len, is_split_into_words=True,
                    return_offsets_mapping=True)
    for i, (_, tags) in enumerate(sentences):
        word_ids = enc.word_ids(batch_index=i)
        labs, prev = [], None
        for wid in word_ids:
            if wid is None:
                labs.append(-100)
            elif wid != prev:
                labs.append(label_to_id[tags[wid]])
            else:
                labs.append(-100)
            prev = wid
        feats.append({"input_ids": enc["input_ids"][i], "labels": labs,
                      "token_type_ids": enc["token_type_ids"][i] if "token_type_ids" in enc else [0] * len(labs),
                      "attention_mask": enc["attention_mask"][i]})
    return feats
