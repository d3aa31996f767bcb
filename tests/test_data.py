"""Datasets and transforms with the reference's call signatures (launch_dist.py:64-74, example_launch.py:32-60)."""
import gzip
import os
import pickle
import struct

import numpy as np
import pytest
import torch

import pytorch_distributed_learning_b200 as pdl
from pytorch_distributed_learning_b200.utils.data import CIFAR10, MNIST, DataLoader, transforms as T
from pytorch_distributed_learning_b200.utils.data.distributed import DistributedSampler


def _write_idx(path, arr, gz=False):
    payload = struct.pack(">HBB", 0, 8, arr.ndim) + struct.pack(">" + "I" * arr.ndim, *arr.shape) + arr.tobytes()
    with (gzip.open(path + ".gz", "wb") if gz else open(path, "wb")) as f:
        f.write(payload)


@pytest.mark.parametrize("gz", [False, True])
def test_mnist_reads_idx_files(tmp_path, gz):
    rng = np.random.default_rng(0)
    images = rng.integers(0, 256, size=(37, 28, 28), dtype=np.uint8)
    labels = rng.integers(0, 10, size=(37,), dtype=np.uint8)
    raw = tmp_path / "MNIST" / "raw"
    raw.mkdir(parents=True)
    _write_idx(str(raw / "train-images-idx3-ubyte"), images, gz)
    _write_idx(str(raw / "train-labels-idx1-ubyte"), labels, gz)
    ds = MNIST(root=str(tmp_path), train=True, transform=T.mnist_transform(), download=True)
    assert not ds.synthetic and len(ds) == 37
    x, y = ds[5]
    assert x.shape == (1, 28, 28) and x.dtype == torch.float32 and y == int(labels[5])
    ref = (torch.from_numpy(images[5]).float() / 255 - 0.1307) / 0.3081
    torch.testing.assert_close(x[0], ref)
    # sharded exactly like the reference: DistributedSampler + DataLoader(shuffle=False, pin_memory irrelevant on CPU)
    sampler = DistributedSampler(ds, num_replicas=2, rank=1)
    xb, yb = next(iter(DataLoader(ds, batch_size=8, shuffle=False, sampler=sampler)))
    assert xb.shape == (8, 1, 28, 28) and yb.dtype == torch.int64


def test_cifar10_reads_pickles(tmp_path):
    rng = np.random.default_rng(1)
    base = tmp_path / "cifar-10-batches-py"
    base.mkdir()
    all_data, all_labels = [], []
    for i in range(1, 6):
        data = rng.integers(0, 256, size=(4, 3072), dtype=np.uint8)
        labels = rng.integers(0, 10, size=(4,)).tolist()
        pickle.dump({"data": data, "labels": labels}, open(base / ("data_batch_%d" % i), "wb"))
        all_data.append(data); all_labels += labels
    ds = CIFAR10(root=str(tmp_path), train=True, download=False, transform=T.Compose([T.ToTensor()]))
    assert not ds.synthetic and len(ds) == 20
    x, y = ds[7]
    assert x.shape == (3, 32, 32) and y == all_labels[7]
    torch.testing.assert_close(x, torch.from_numpy(np.concatenate(all_data)[7].reshape(3, 32, 32)).float() / 255)


def test_missing_files_fall_back_to_synthetic_or_raise(tmp_path):
    ds = CIFAR10(root=str(tmp_path), train=False, synthetic_length=16, transform=T.cifar10_train_transform())
    assert ds.synthetic and len(ds) == 16
    x, y = ds[3]
    assert x.shape == (3, 32, 32) and 0 <= y < 10
    again = CIFAR10(root=str(tmp_path), train=False, synthetic_length=16)
    assert np.array_equal(ds.data, again.data)  # deterministic
    with pytest.raises(RuntimeError):
        MNIST(root=str(tmp_path), allow_synthetic=False)


def test_transforms_match_their_definitions():
    g = torch.Generator().manual_seed(0)
    img = torch.randint(0, 256, (32, 32, 3), dtype=torch.uint8)
    t = T.ToTensor()(img.numpy())
    torch.testing.assert_close(t, img.permute(2, 0, 1).float() / 255)
    n = T.Normalize(T.CIFAR10_MEAN, T.CIFAR10_STD)(t)
    torch.testing.assert_close(n[1], (t[1] - 0.4822) / 0.1994)
    crop = T.RandomCrop(32, padding=4, generator=g)
    out = crop(img)
    assert out.shape == (32, 32, 3)
    padded = torch.nn.functional.pad(img.permute(2, 0, 1), (4, 4, 4, 4))
    # the crop is some window of the zero-padded image
    assert any(torch.equal(out.permute(2, 0, 1), padded[:, i:i + 32, j:j + 32]) for i in range(9) for j in range(9))
    flip = T.RandomHorizontalFlip(p=1.0)
    assert torch.equal(flip(img), img.flip(1))
    assert torch.equal(T.RandomHorizontalFlip(p=0.0)(img), img)
    assert torch.equal(flip(t), t.flip(-1))  # CxHxW tensors flip along W too
    with pytest.raises(ValueError):
        T.RandomCrop(64)(img)
    assert "RandomCrop" in repr(T.cifar10_train_transform())
    assert pdl.utils.data.transforms is T


def test_cuda_prefetcher_is_a_passthrough_on_cpu():
    import torch

    import pytorch_distributed_learning_b200 as pdl

    ds = pdl.utils.data.SyntheticMNIST(20)
    loader = torch.utils.data.DataLoader(ds, batch_size=6, shuffle=False)
    pre = pdl.utils.data.CudaPrefetcher(loader, "cpu")
    assert len(pre) == len(loader) == 4 and pre.sampler is loader.sampler
    for (a, b), (c, d) in zip(pre, loader):
        assert torch.equal(a, c) and torch.equal(b, d)
    assert sum(1 for _ in pre) == 4  # re-iterable
