#!/usr/bin/env python3
"""Evaluate an MNIST checkpoint written by ``train.py --task mnist`` (API of hetseq/eval_mnist.py:39-100:
``--model_ckpt``, ``--mnist_dir``); prints average loss and accuracy on the test split."""
from __future__ import annotations

import argparse
import os
import sys

if __name__ == "__main__" and __package__ is None:
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    __package__ = "hetseq_9cme_b200"

import torch
import torch.nn.functional as F

from .data import MNISTDataset
from .tasks import MNISTNet


def evaluate(model_ckpt: str, mnist_dir: str, batch_size: int = 1000, device=None):
    device = device or ("cuda" if torch.cuda.is_available() else "cpu")
    model = MNISTNet()
    state = torch.load(model_ckpt, map_location="cpu", weights_only=False)
    model.load_state_dict(state["model"])
    model.to(device).eval()
    proc = os.path.join(mnist_dir, "MNIST", "processed")
    path = os.path.join(proc if os.path.isdir(proc) else mnist_dir, "test.pt")
    ds = MNISTDataset(path)
    loss_sum, correct = 0.0, 0
    with torch.no_grad():
        for s in range(0, len(ds), batch_size):
            x, y = ds.fetch_batch(range(s, min(len(ds), s + batch_size)))
            x, y = x.to(device), y.to(device)
            out = model(x, y, eval=True)
            loss_sum += F.nll_loss(out, y, reduction="sum").item()
            correct += (out.argmax(1) == y).sum().item()
    n = len(ds)
    print("\nTest set: Average loss: {:.4f}, Accuracy: {}/{} ({:.0f}%)\n".format(loss_sum / n, correct, n, 100.0 * correct / n))
    return loss_sum / n, correct / n


def main(argv=None):
    ap = argparse.ArgumentParser()
    ap.add_argument("--model_ckpt", required=True)
    ap.add_argument("--mnist_dir", required=True)
    a = ap.parse_args(argv)
    evaluate(a.model_ckpt, a.mnist_dir)


if __name__ == "__main__":
    main()
