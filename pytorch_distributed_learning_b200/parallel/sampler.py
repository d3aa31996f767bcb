"""DistributedSampler: shards a dataset's indices over the ranks of the default group.

Index streams are bit-identical to the sampler the reference uses (launch_dist.py:67-71,
example_launch.py:50-53): ``num_samples = ceil(len / replicas)``, optional shuffle with a
``randperm`` seeded by ``seed + epoch``, pad by repeating the head of the list (or drop the
tail), then take ``indices[rank::replicas]``.  ``set_epoch`` re-seeds the shuffle
(example_mp.py:100); scripts that never call it see the same order every epoch.
"""
from __future__ import annotations

import math

import torch
from torch.utils.data import Sampler

from . import distributed as dist


class DistributedSampler(Sampler):
    def __init__(self, dataset, num_replicas=None, rank=None, shuffle=True, seed=0, drop_last=False):
        if num_replicas is None:
            if not dist.is_initialized():
                raise RuntimeError("Requires distributed package to be available")
            num_replicas = dist.get_world_size()
        if rank is None:
            if not dist.is_initialized():
                raise RuntimeError("Requires distributed package to be available")
            rank = dist.get_rank()
        if rank >= num_replicas or rank < 0:
            raise ValueError("Invalid rank %d, rank should be in the interval [0, %d]" % (rank, num_replicas - 1))
        self.dataset, self.num_replicas, self.rank = dataset, num_replicas, rank
        self.epoch, self.drop_last, self.shuffle, self.seed = 0, drop_last, shuffle, seed
        n = len(dataset)
        if drop_last and n % num_replicas != 0:
            self.num_samples = math.ceil((n - num_replicas) / num_replicas)
        else:
            self.num_samples = math.ceil(n / num_replicas)
        self.total_size = self.num_samples * num_replicas

    def __iter__(self):
        n = len(self.dataset)
        if self.shuffle:
            g = torch.Generator()
            g.manual_seed(self.seed + self.epoch)
            indices = torch.randperm(n, generator=g).tolist()
        else:
            indices = list(range(n))
        if not self.drop_last:
            pad = self.total_size - len(indices)
            if pad <= len(indices):
                indices += indices[:pad]
            else:
                indices += (indices * math.ceil(pad / len(indices)))[:pad]
        else:
            indices = indices[: self.total_size]
        assert len(indices) == self.total_size
        indices = indices[self.rank : self.total_size : self.num_replicas]
        assert len(indices) == self.num_samples
        return iter(indices)

    def __len__(self):
        return self.num_samples

    def set_epoch(self, epoch):
        self.epoch = epoch
