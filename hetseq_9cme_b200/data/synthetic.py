"""Synthetic corpora with the shapes of the real ones (there is no network / no datasets on the box).

* ``make_bert_shards``  — NVIDIA-style pre-training shards (six int arrays) as ``.b2s`` or ``.hdf5``
  (the latter through the built-in minimal HDF5 writer);
* ``make_mnist``        — ``MNIST/processed/{training,test}.pt`` tuples of (uint8 images, labels) with a learnable
  class-dependent pattern so that a few hundred steps visibly reduce the loss;
* ``make_conll``        — CoNLL-2003-shaped token/tag files + vocabulary for the NER task.
"""
from __future__ import annotations

import json
import os
from typing import List, Optional

import numpy as np
import torch


def bert_shard_arrays(n: int, seq_len: int = 128, max_pred: int = 20, vocab_size: int = 30522, seed: int = 0,
                      full_length: bool = True):
    rng = np.random.RandomState(seed)
    low = min(1000, max(1, vocab_size // 2))
    ids = rng.randint(low, vocab_size, size=(n, seq_len)).astype(np.int32)
    ids[:, 0] = min(101, vocab_size - 1)
    lens = np.full(n, seq_len) if full_length else rng.randint(seq_len // 2, seq_len + 1, size=n)
    mask = (np.arange(seq_len)[None, :] < lens[:, None]).astype(np.int32)
    ids *= mask
    first = (lens // 2)[:, None]
    seg = ((np.arange(seq_len)[None, :] >= first) & (mask > 0)).astype(np.int32)
    pos = np.zeros((n, max_pred), dtype=np.int32)
    mids = np.zeros((n, max_pred), dtype=np.int32)
    for r in range(n):
        k = rng.randint(max(1, max_pred - 3), max_pred + 1)
        p = np.sort(rng.choice(np.arange(1, lens[r]), size=min(k, lens[r] - 1), replace=False))
        pos[r, :len(p)] = p
        mids[r, :len(p)] = rng.randint(low, vocab_size, size=len(p))
    nsl = rng.randint(0, 2, size=n).astype(np.int32)
    return {"input_ids": ids, "input_mask": mask, "segment_ids": seg, "masked_lm_positions": pos,
            "masked_lm_ids": mids, "next_sentence_labels": nsl}


def make_bert_shards(root: str, n_train_files: int = 2, samples_per_file: int = 256, seq_len: int = 128, max_pred: int = 20,
                     vocab_size: int = 30522, fmt: str = "b2s", seed: int = 0, test_samples: int = 64) -> str:
    from . import shard_io
    os.makedirs(root, exist_ok=True)
    def write(name, arrays):
        path = os.path.join(root, name + ("." + fmt))
        if fmt == "b2s":
            shard_io.write_b2shard(path, arrays)
        elif fmt in ("hdf5", "h5"):
            from .hdf5_lite import write_hdf5
            write_hdf5(path, arrays)
        else:
            raise ValueError(fmt)
    for i in range(n_train_files):
        write("train_shard_%d" % i, bert_shard_arrays(samples_per_file, seq_len, max_pred, vocab_size, seed + i))
    write("test_shard_0", bert_shard_arrays(test_samples, seq_len, max_pred, vocab_size, seed + 1000))
    return root


def write_bert_config(path: str, **kw) -> str:
    cfg = dict(vocab_size=30522, hidden_size=768, num_hidden_layers=12, num_attention_heads=12, intermediate_size=3072,
               hidden_act="gelu", hidden_dropout_prob=0.1, attention_probs_dropout_prob=0.1, max_position_embeddings=512,
               type_vocab_size=2, initializer_range=0.02)
    cfg.update(kw)
    with open(path, "w") as f:
        json.dump(cfg, f, indent=2)
    return path


def write_vocab(path: str, size: int = 30522) -> str:
    specials = ["[PAD]"] + ["[unused%d]" % i for i in range(99)] + ["[UNK]", "[CLS]", "[SEP]", "[MASK]"]
    with open(path, "w", encoding="utf-8") as f:
        for i in range(size):
            f.write((specials[i] if i < len(specials) else "tok%d" % i) + "\n")
    return path


def make_mnist(root: str, n_train: int = 2048, n_test: int = 512, seed: int = 0) -> str:
    """Images = class-dependent blob + noise, written in torchvision's processed layout."""
    rng = np.random.RandomState(seed)
    proc = os.path.join(root, "MNIST", "processed")
    os.makedirs(proc, exist_ok=True)
    yy, xx = np.mgrid[0:28, 0:28]

    def gen(n):
        labels = rng.randint(0, 10, size=n)
        cx = 6 + 2 * (labels % 5) * 2
        cy = 8 + 12 * (labels // 5)
        img = np.exp(-(((xx[None] - cx[:, None, None]) ** 2 + (yy[None] - cy[:, None, None]) ** 2) / 18.0)) * 220
        img = img + rng.rand(n, 28, 28) * 35
        return torch.from_numpy(img.clip(0, 255).astype(np.uint8)), torch.from_numpy(labels.astype(np.int64))

    torch.save(gen(n_train), os.path.join(proc, "training.pt"))
    torch.save(gen(n_test), os.path.join(proc, "test.pt"))
    return root


CONLL_TAGS = ["O", "B-PER", "I-PER", "B-ORG", "I-ORG", "B-LOC", "I-LOC", "B-MISC", "I-MISC"]


def make_conll(root: str, n_train: int = 256, n_valid: int = 64, n_test: int = 64, vocab_words: int = 500, seed: int = 0,
               entities: int = 0, entity_dim: int = 16):
    """CoNLL-2003-shaped files: ``token POS chunk NER`` per line, blank line between sentences; plus vocab.txt.
    ``entities > 0`` adds what the entity-linking task reads: a fifth column with the linked entity's name (``O`` when
    unlinked), ``entity_vocab.txt`` (line number = id, id 0 = no entity) and ``ent_vecs.pt`` ([entities + 1, entity_dim])."""
    rng = np.random.RandomState(seed)
    os.makedirs(root, exist_ok=True)
    words = ["w%d" % i for i in range(vocab_words)]

    def sentence():
        n = rng.randint(4, 40)
        toks, tags, ents = [], [], []
        i = 0
        while i < n:
            if rng.rand() < 0.15:
                kind = rng.choice(["PER", "ORG", "LOC", "MISC"])
                ln = rng.randint(1, 4)
                ent = "E%d" % rng.randint(1, entities + 1) if entities else "O"
                for j in range(ln):
                    toks.append(words[rng.randint(vocab_words // 2, vocab_words)])
                    tags.append(("B-" if j == 0 else "I-") + kind)
                    ents.append(ent)
                i += ln
            else:
                toks.append(words[rng.randint(0, vocab_words // 2)])
                tags.append("O")
                ents.append("O")
                i += 1
        return toks, tags, ents

    def write(name, n):
        path = os.path.join(root, name)
        with open(path, "w") as f:
            f.write("-DOCSTART- -X- -X- O\n\n")
            for _ in range(n):
                toks, tags, ents = sentence()
                for t, g, e in zip(toks, tags, ents):
                    f.write("%s NN B-NP %s%s\n" % (t, g, (" " + e) if entities else ""))
                f.write("\n")
        return path

    paths = {"train": write("train.txt", n_train), "validation": write("valid.txt", n_valid), "test": write("test.txt", n_test)}
    vocab = os.path.join(root, "vocab.txt")
    with open(vocab, "w") as f:
        for t in ["[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]"] + words:
            f.write(t + "\n")
    paths["vocab"] = vocab
    if entities:
        import torch
        with open(os.path.join(root, "entity_vocab.txt"), "w") as f:
            for name in ["O"] + ["E%d" % k for k in range(1, entities + 1)]:
                f.write(name + "\n")
        vecs = torch.from_numpy(rng.randn(entities + 1, entity_dim).astype(np.float32))
        paths["ent_vecs"] = os.path.join(root, "ent_vecs.pt")
        torch.save(vecs, paths["ent_vecs"])
    return paths
