"""The MNIST-shaped ConvNet of the reference (launch_dist.py:9-41, mpspawn_dist.py:11-43).

Layers, hyper-parameters, attribute names and therefore ``state_dict`` keys are the
reference's: conv1 1->32 k5 p1, pool 2/2, conv2 32->64 k3, pool 2/1 (overlapping), conv3
64->128 k3, pool 2/2, flatten 2048, fc1 2048->10.  ``dropout`` exists but is not applied in
``forward`` -- exactly like the reference.  Built from this package's ``nn`` layers, so on a
B200 every op is a first-party kernel; ``TrainStep`` additionally recognises this class and
runs its fused conv+bias+ReLU+pool pipeline inside one CUDA graph.
"""
from __future__ import annotations

import torch

from .. import nn

_CHANNELS_LAST = torch.channels_last


class ConvNet(nn.Module):
    def __init__(self):
        super().__init__()
        self.relu = nn.ReLU()
        self.conv1 = nn.Conv2d(in_channels=1, out_channels=32, kernel_size=5, stride=1, padding=1)
        self.maxpool1 = nn.MaxPool2d(kernel_size=2, stride=2)
        self.conv2 = nn.Conv2d(in_channels=32, out_channels=64, kernel_size=3, stride=1)
        self.maxpool2 = nn.MaxPool2d(kernel_size=2, stride=1)
        self.conv3 = nn.Conv2d(in_channels=64, out_channels=128, kernel_size=3, stride=1)
        self.maxpool3 = nn.MaxPool2d(kernel_size=2, stride=2)
        self.dropout = nn.Dropout(p=0.5)
        self.fc1 = nn.Linear(in_features=128 * 4 * 4, out_features=10)

    def _block(self, conv, pool, x):
        """``pool(relu(conv(x)))``: one fused kernel on a B200 (``ops.functional.conv_relu_pool``), the three layers otherwise."""
        if x.is_cuda and x.dtype == torch.float32 and type(conv) is nn.Conv2d and type(pool) is nn.MaxPool2d:
            from ..ops import functional as NF

            y = NF.conv_relu_pool(x, conv, pool)
            if y is not None:
                return y
        return pool(self.relu(conv(x)))

    def forward(self, x):
        x = self._block(self.conv1, self.maxpool1, x)
        x = self._block(self.conv2, self.maxpool2, x)
        x = self._block(self.conv3, self.maxpool3, x)
        if x.is_cuda and x.dim() == 4 and x.is_contiguous(memory_format=_CHANNELS_LAST) and not x.is_contiguous():
            from ..ops import functional as NF

            x = NF.flatten_nchw(x)  # the reference's reshape(-1, 2048) in NCHW element order, as one permute kernel
        else:
            x = x.reshape(-1, 128 * 4 * 4)
        return self.fc1(x)
