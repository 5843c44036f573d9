"""Synthetic datasets of the shapes the benchmarks name (no network, no dataset files).

``SyntheticImages`` yields CIFAR-10-shaped (3x32x32, 10 classes) or ImageNet-shaped (3x224x224, 1000 classes)
fp32 batches; ``SyntheticTokens`` yields BERT-shaped token batches.  Batches live in pinned host memory so the
training step's host->device copy is a true asynchronous DMA (the reference's loaders are not pinned,
/root/reference/code/distributed_training/model_parallel.py:91-92, which makes its ``non_blocking=True`` moot).
"""
from __future__ import annotations

import torch
from torch.utils.data import Dataset


class SyntheticImages(Dataset):
    def __init__(self, length=50000, shape=(3, 32, 32), num_classes=10, seed=0, pool=64):
        self.length, self.shape, self.num_classes = length, tuple(shape), num_classes
        g = torch.Generator().manual_seed(seed)
        self.pool = torch.randn((pool,) + self.shape, generator=g)
        # the label is a function of the image (one class per pool entry), so the set is learnable: trainers can show a
        # rising accuracy offline, the way the reference's curves do on CIFAR-10 (Readme.md:283-294)
        pool_labels = torch.randint(0, num_classes, (pool,), generator=g)
        self.labels = pool_labels[torch.arange(length) % pool]

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        return self.pool[i % self.pool.shape[0]], int(self.labels[i])


class SyntheticTokens(Dataset):
    def __init__(self, length=10000, seq_len=128, vocab=30522, num_classes=2, seed=0):
        g = torch.Generator().manual_seed(seed)
        self.ids = torch.randint(0, vocab, (min(length, 512), seq_len), generator=g)
        self.labels = torch.randint(0, num_classes, (length,), generator=g)
        self.length = length

    def __len__(self):
        return self.length

    def __getitem__(self, i):
        return self.ids[i % self.ids.shape[0]], int(self.labels[i])


def pinned_batches(n_batches: int, batch: int, shape=(3, 32, 32), num_classes=10, seed=0, tokens=False, vocab=30522):
    """``n_batches`` distinct (inputs, targets) pairs in pinned host memory."""
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(n_batches):
        if tokens:
            x = torch.randint(0, vocab, (batch,) + tuple(shape), generator=g)
        else:
            x = torch.randn((batch,) + tuple(shape), generator=g)
        y = torch.randint(0, num_classes, (batch,), generator=g)
        if torch.cuda.is_available():
            x, y = x.pin_memory(), y.pin_memory()
        out.append((x, y))
    return out
