"""``MNIST`` and ``CIFAR10`` with the constructor the reference uses (torchvision's), no torchvision required.

Reference call sites: ``torchvision.datasets.MNIST(root='./data', train=True, transform=..., download=True)``
(launch_dist.py:65, mpspawn_dist.py:73) and ``torchvision.datasets.CIFAR10(root=..., train=True, download=False,
transform=...)`` (example_launch.py:32-48, example_mp.py:56-72).

The files are read in their published on-disk formats when they are present under ``root`` (MNIST ``*-idx3-ubyte`` /
``*-idx1-ubyte``, optionally gzipped, in ``root``, ``root/MNIST/raw`` or ``root/raw``; CIFAR-10 ``cifar-10-batches-py``
pickles).  The training boxes have no network, so ``download=True`` cannot fetch anything: when the files are absent
the dataset falls back to deterministic synthetic ``uint8`` images of the right shape (``synthetic=True`` is then set
on the object and a one-line notice is printed on rank 0) unless ``allow_synthetic=False``, which raises instead.
Samples go through ``transform`` / ``target_transform`` exactly like torchvision's: ``(image, int label)``.
"""
from __future__ import annotations

import gzip
import os
import pickle
import struct
import sys

import numpy as np
import torch
from torch.utils.data import Dataset


def _notice(msg):
    if os.environ.get("RANK", "0") == "0":
        sys.stderr.write("[pdl.data] %s\n" % msg)


def _open_maybe_gz(path):
    if os.path.exists(path):
        return open(path, "rb")
    if os.path.exists(path + ".gz"):
        return gzip.open(path + ".gz", "rb")
    return None


def read_idx(path):
    """One IDX file (the MNIST container): magic 0x0000 | dtype 0x08 | ndim, big-endian dims, raw uint8 payload."""
    f = _open_maybe_gz(path)
    if f is None:
        return None
    with f:
        zero, dtype, ndim = struct.unpack(">HBB", f.read(4))
        if zero != 0 or dtype != 0x08:
            raise ValueError("%s is not a uint8 IDX file (magic %04x %02x)" % (path, zero, dtype))
        dims = struct.unpack(">" + "I" * ndim, f.read(4 * ndim))
        data = np.frombuffer(f.read(), dtype=np.uint8)
    if data.size != int(np.prod(dims)):
        raise ValueError("%s: payload of %d bytes does not match dims %s" % (path, data.size, dims))
    return data.reshape(dims)


class _VisionDataset(Dataset):
    shape = None  # H, W[, C] of one uint8 sample
    classes = 10

    def __init__(self, root, train, transform, target_transform, download, allow_synthetic, synthetic_length):
        self.root, self.train = root, bool(train)
        self.transform, self.target_transform = transform, target_transform
        loaded = self._load()
        self.synthetic = loaded is None
        if loaded is None:
            if not allow_synthetic:
                raise RuntimeError("%s files not found under %r and there is no network to download them" % (type(self).__name__, root))
            n = synthetic_length or self.default_length[0 if self.train else 1]
            g = np.random.default_rng(1234 if self.train else 4321)
            loaded = (g.integers(0, 256, size=(n,) + tuple(self.shape), dtype=np.uint8), g.integers(0, self.classes, size=(n,), dtype=np.int64))
            _notice("%s: no files under %r%s; using %d deterministic synthetic samples of shape %s" % (
                type(self).__name__, root, " (download=True ignored: no network)" if download else "", n, self.shape))
        self.data, self.targets = loaded

    def __len__(self):
        return len(self.targets)

    def __getitem__(self, i):
        img, target = self.data[i], int(self.targets[i])
        if self.transform is not None:
            img = self.transform(img)
        if self.target_transform is not None:
            target = self.target_transform(target)
        return img, target


class MNIST(_VisionDataset):
    shape = (28, 28)
    default_length = (60000, 10000)

    def __init__(self, root="./data", train=True, transform=None, target_transform=None, download=False, allow_synthetic=True,
                 synthetic_length=None):
        super().__init__(root, train, transform, target_transform, download, allow_synthetic, synthetic_length)

    def _load(self):
        stem = "train" if self.train else "t10k"
        for sub in ("", os.path.join("MNIST", "raw"), "raw"):
            base = os.path.join(self.root, sub)
            images = read_idx(os.path.join(base, "%s-images-idx3-ubyte" % stem))
            labels = read_idx(os.path.join(base, "%s-labels-idx1-ubyte" % stem))
            if images is not None and labels is not None:
                if images.shape[0] != labels.shape[0] or images.shape[1:] != (28, 28):
                    raise ValueError("MNIST files under %r are inconsistent: %s images, %s labels" % (base, images.shape, labels.shape))
                return images, labels.astype(np.int64)
        return None


class CIFAR10(_VisionDataset):
    shape = (32, 32, 3)
    default_length = (50000, 10000)

    def __init__(self, root="./data", train=True, transform=None, target_transform=None, download=False, allow_synthetic=True,
                 synthetic_length=None):
        super().__init__(root, train, transform, target_transform, download, allow_synthetic, synthetic_length)

    def _load(self):
        base = os.path.join(self.root, "cifar-10-batches-py")
        names = ["data_batch_%d" % i for i in range(1, 6)] if self.train else ["test_batch"]
        if not all(os.path.exists(os.path.join(base, n)) for n in names):
            return None
        data, labels = [], []
        for n in names:
            with open(os.path.join(base, n), "rb") as f:
                entry = pickle.load(f, encoding="latin1")
            data.append(np.asarray(entry["data"], dtype=np.uint8))
            labels.extend(entry.get("labels", entry.get("fine_labels")))
        images = np.concatenate(data).reshape(-1, 3, 32, 32).transpose(0, 2, 3, 1)  # stored CHW planes -> HWC like torchvision
        return np.ascontiguousarray(images), np.asarray(labels, dtype=np.int64)


def collate_to_tensors(batch):
    """Default-collate equivalent for (tensor image, int label) samples."""
    return torch.stack([b[0] for b in batch]), torch.tensor([b[1] for b in batch], dtype=torch.int64)
