import itertools

import pytest
import torch
from torch.utils.data.distributed import DistributedSampler as TorchSampler

from pytorch_distributed_learning_b200.utils.data.distributed import DistributedSampler


class _DS:
    def __init__(self, n):
        self.n = n

    def __len__(self):
        return self.n


@pytest.mark.parametrize("n,world,shuffle,drop_last", list(itertools.product([1, 7, 100, 1001], [1, 2, 3, 8], [False, True], [False, True])))
def test_matches_torch(n, world, shuffle, drop_last):
    ds = _DS(n)
    for rank in range(world):
        ours = DistributedSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=5, drop_last=drop_last)
        ref = TorchSampler(ds, num_replicas=world, rank=rank, shuffle=shuffle, seed=5, drop_last=drop_last)
        for epoch in (0, 1, 4):
            ours.set_epoch(epoch)
            ref.set_epoch(epoch)
            assert list(ours) == list(ref)
            assert len(ours) == len(ref)


def test_mnist_step_count():
    # launch_dist.py:77 prints len(train_loader): 60000 samples, W ranks, batch 100
    for world in (1, 2, 8):
        s = DistributedSampler(_DS(60000), num_replicas=world, rank=0)
        assert -(-len(s) // 100) == -(-(-(-60000 // world)) // 100)


def test_no_set_epoch_repeats_order():
    s = DistributedSampler(_DS(50), num_replicas=2, rank=1)
    assert list(s) == list(s)  # ConvNet scripts never call set_epoch
    s.set_epoch(1)
    first = list(s)
    s.set_epoch(2)
    assert first != list(s)


def test_invalid_rank():
    with pytest.raises(ValueError):
        DistributedSampler(_DS(4), num_replicas=2, rank=2)
    with pytest.raises(RuntimeError):
        DistributedSampler(_DS(4))  # no default group
