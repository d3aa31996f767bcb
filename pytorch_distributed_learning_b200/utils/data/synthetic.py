"""Synthetic stand-ins for the datasets the reference downloads (MNIST: launch_dist.py:65,
CIFAR-10: example_launch.py:32).  There is no network on the training boxes, so the examples
and benchmarks draw deterministic random tensors of the same shape/normalisation instead."""
from __future__ import annotations

import torch
from torch.utils.data import Dataset


class SyntheticDataset(Dataset):
    """``length`` samples of shape ``shape`` ~ N(0,1) and labels in [0, num_classes).

    Samples are generated once (seeded) and indexed, so a DistributedSampler shards them the
    same way it would shard a real dataset."""

    def __init__(self, length, shape, num_classes=10, seed=0, dtype=torch.float32):
        g = torch.Generator().manual_seed(seed)
        self.images = torch.randn((length,) + tuple(shape), generator=g, dtype=dtype)
        self.labels = torch.randint(0, num_classes, (length,), generator=g, dtype=torch.int64)

    def __len__(self):
        return self.images.shape[0]

    def __getitem__(self, i):
        return self.images[i], self.labels[i]


class SyntheticMNIST(SyntheticDataset):
    def __init__(self, length=60000, seed=0):
        super().__init__(length, (1, 28, 28), 10, seed)


class SyntheticCIFAR10(SyntheticDataset):
    def __init__(self, length=50000, seed=0):
        super().__init__(length, (3, 32, 32), 10, seed)


class SyntheticImageNet(SyntheticDataset):
    def __init__(self, length=1024, seed=0, num_classes=1000):
        super().__init__(length, (3, 224, 224), num_classes, seed)
