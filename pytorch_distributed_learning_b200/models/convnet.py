"""The MNIST-shaped ConvNet of the reference (launch_dist.py:9-41, mpspawn_dist.py:11-43).

Layers, hyper-parameters, attribute names and therefore ``state_dict`` keys are the
reference's: conv1 1->32 k5 p1, pool 2/2, conv2 32->64 k3, pool 2/1 (overlapping), conv3
64->128 k3, pool 2/2, flatten 2048, fc1 2048->10.  ``dropout`` exists but is not applied in
``forward`` -- exactly like the reference.  Built from this package's ``nn`` layers, so on a
B200 every op is a first-party kernel; ``TrainStep`` additionally recognises this class and
runs its fused conv+bias+ReLU+pool pipeline inside one CUDA graph.
"""
from __future__ import annotations

from .. import nn


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

    def forward(self, x):
        x = self.maxpool1(self.relu(self.conv1(x)))
        x = self.maxpool2(self.relu(self.conv2(x)))
        x = self.maxpool3(self.relu(self.conv3(x)))
        x = x.reshape(-1, 128 * 4 * 4)
        return self.fc1(x)
