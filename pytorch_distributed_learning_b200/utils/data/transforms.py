"""The input transforms the reference's scripts compose (tensor-native, no PIL / torchvision needed).

Reference call sites: MNIST ``ToTensor`` + ``Normalize((0.1307,), (0.3081,))`` (launch_dist.py:64-65); CIFAR-10
``RandomCrop(32, padding=4)`` + ``RandomHorizontalFlip`` + ``ToTensor`` + ``Normalize(mean, std)``
(example_launch.py:32-46).  Each transform maps one sample: ``uint8 HxWxC / HxW`` arrays or ``CxHxW`` tensors in,
``CxHxW`` float tensors out, so they plug into ``torch.utils.data.Dataset.__getitem__`` exactly like torchvision's.
"""
from __future__ import annotations

import numpy as np
import torch

MNIST_MEAN, MNIST_STD = (0.1307,), (0.3081,)
CIFAR10_MEAN, CIFAR10_STD = (0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010)


class Compose:
    def __init__(self, transforms):
        self.transforms = list(transforms)

    def __call__(self, x):
        for t in self.transforms:
            x = t(x)
        return x

    def __repr__(self):
        return "Compose(%s)" % ", ".join(repr(t) for t in self.transforms)


class ToTensor:
    """uint8 ``HxWxC`` / ``HxW`` (numpy or tensor) -> float32 ``CxHxW`` in [0, 1]; float inputs only get the layout change."""

    def __call__(self, x):
        t = torch.from_numpy(np.ascontiguousarray(x)) if isinstance(x, np.ndarray) else x
        if not torch.is_tensor(t):
            raise TypeError("ToTensor expects a numpy array or a tensor, got %s" % type(x).__name__)
        if t.dim() == 2:
            t = t.unsqueeze(-1)
        if t.dim() != 3:
            raise ValueError("ToTensor expects HxW or HxWxC, got shape %s" % (tuple(t.shape),))
        t = t.permute(2, 0, 1).contiguous()
        return t.to(torch.float32).div_(255.0) if t.dtype == torch.uint8 else t.to(torch.float32)

    def __repr__(self):
        return "ToTensor()"


class Normalize:
    def __init__(self, mean, std):
        self.mean = torch.as_tensor(mean, dtype=torch.float32).view(-1, 1, 1)
        self.std = torch.as_tensor(std, dtype=torch.float32).view(-1, 1, 1)
        if (self.std == 0).any():
            raise ValueError("std must be non-zero")

    def __call__(self, x):
        return (x - self.mean) / self.std

    def __repr__(self):
        return "Normalize(mean=%s, std=%s)" % (self.mean.flatten().tolist(), self.std.flatten().tolist())


def _chw(x):
    """Samples may reach the random transforms before ToTensor (torchvision order): view HxWxC as CxHxW and back."""
    if isinstance(x, np.ndarray):
        x = torch.from_numpy(np.ascontiguousarray(x))
    hwc = x.dim() == 3 and x.shape[-1] <= 4 < x.shape[0]
    if x.dim() == 2:
        return x.unsqueeze(0), "hw"
    return (x.permute(2, 0, 1), "hwc") if hwc else (x, "chw")


def _restore(x, kind):
    return x.squeeze(0) if kind == "hw" else (x.permute(1, 2, 0) if kind == "hwc" else x)


class RandomCrop:
    """Zero-pad by ``padding`` on every side, then take a uniformly placed ``size`` x ``size`` window."""

    def __init__(self, size, padding=0, generator=None):
        self.size = (size, size) if isinstance(size, int) else tuple(size)
        self.padding = int(padding)
        self.generator = generator

    def __call__(self, x):
        t, kind = _chw(x)
        if self.padding:
            t = torch.nn.functional.pad(t, (self.padding,) * 4)
        h, w = t.shape[-2:]
        th, tw = self.size
        if th > h or tw > w:
            raise ValueError("crop %s is larger than the padded input %s" % (self.size, (h, w)))
        i = int(torch.randint(0, h - th + 1, (1,), generator=self.generator))
        j = int(torch.randint(0, w - tw + 1, (1,), generator=self.generator))
        return _restore(t[..., i:i + th, j:j + tw], kind)

    def __repr__(self):
        return "RandomCrop(size=%s, padding=%d)" % (self.size, self.padding)


class RandomHorizontalFlip:
    def __init__(self, p=0.5, generator=None):
        self.p = float(p)
        self.generator = generator

    def __call__(self, x):
        if float(torch.rand(1, generator=self.generator)) >= self.p:
            return x
        t, kind = _chw(x)
        return _restore(t.flip(-1), kind)

    def __repr__(self):
        return "RandomHorizontalFlip(p=%g)" % self.p


def mnist_transform():
    """launch_dist.py:64-65."""
    return Compose([ToTensor(), Normalize(MNIST_MEAN, MNIST_STD)])


def cifar10_train_transform():
    """example_launch.py:32-46."""
    return Compose([RandomCrop(32, padding=4), RandomHorizontalFlip(), ToTensor(), Normalize(CIFAR10_MEAN, CIFAR10_STD)])
