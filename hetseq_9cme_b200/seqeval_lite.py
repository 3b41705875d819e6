"""Entity-level sequence-labelling metrics (what the reference obtains from the ``seqeval`` package in
test/test_eval_bert_fine_tuning.py:130-169; seqeval is not installed here).  IOB1/IOB2 chunks, micro-averaged
precision / recall / F1 over exact-match entities, plus token accuracy."""
from __future__ import annotations

from typing import List, Sequence, Set, Tuple


def get_entities(seq: Sequence[str]) -> Set[Tuple[str, int, int]]:
    """(type, start, end) chunks of one tag sequence (seqeval's default, lenient IOB handling)."""
    chunks = set()
    prev_tag, prev_type, start = "O", "", 0
    for i, label in enumerate(list(seq) + ["O"]):
        tag, typ = (label.split("-", 1) + [""])[:2] if label != "O" else ("O", "")
        end_prev = prev_tag != "O" and (tag in ("B", "O") or typ != prev_type)
        if end_prev:
            chunks.add((prev_type, start, i - 1))
        if tag != "O" and (prev_tag == "O" or tag == "B" or typ != prev_type):
            start = i
        prev_tag, prev_type = tag, typ
    return chunks


def _counts(y_true: List[List[str]], y_pred: List[List[str]]):
    tp = n_pred = n_true = 0
    for t, p in zip(y_true, y_pred):
        et, ep = get_entities(t), get_entities(p)
        tp += len(et & ep)
        n_true += len(et)
        n_pred += len(ep)
    return tp, n_pred, n_true


def precision_score(y_true, y_pred) -> float:
    tp, n_pred, _ = _counts(y_true, y_pred)
    return tp / n_pred if n_pred else 0.0


def recall_score(y_true, y_pred) -> float:
    tp, _, n_true = _counts(y_true, y_pred)
    return tp / n_true if n_true else 0.0


def f1_score(y_true, y_pred) -> float:
    p, r = precision_score(y_true, y_pred), recall_score(y_true, y_pred)
    return 2 * p * r / (p + r) if p + r else 0.0


def accuracy_score(y_true, y_pred) -> float:
    n = sum(len(t) for t in y_true)
    ok = sum(a == b for t, p in zip(y_true, y_pred) for a, b in zip(t, p))
    return ok / n if n else 0.0
