"""ResNet-18/34/50 with torchvision-compatible structure and ``state_dict`` keys.

The reference trains ``torchvision.models.resnet18(pretrained=False, num_classes=10)``
(example_launch.py:26, example_mp.py:50); BASELINE.json additionally names ResNet-50 at
224x224.  These definitions use this package's ``nn`` layers so conv / batch-norm / pooling /
linear run on first-party kernels, and a checkpoint moves between the two implementations
unchanged (same parameter and buffer names: ``conv1.weight``, ``layer1.0.bn1.running_mean``,
``fc.bias`` ...).  Initialisation follows the same scheme (kaiming-normal fan_out for convs,
ones/zeros for batch-norm).
"""
from __future__ import annotations

import os

from .. import nn


def _conv3x3(cin, cout, stride=1):
    return nn.Conv2d(cin, cout, kernel_size=3, stride=stride, padding=1, bias=False)


def _conv1x1(cin, cout, stride=1):
    return nn.Conv2d(cin, cout, kernel_size=1, stride=stride, bias=False)


def _conv_bn(conv, bn, x, relu=False, residual=None, dual=False):
    """conv -> batch norm (-> + residual) (-> ReLU) as the fused kernels see it: the convolution's epilogue reduces the
    normalisation statistics (``bn_stats=True`` is a hint, ignored where it does not apply), and the normalisation,
    skip connection and activation are one pass over the tensor.  ``dual``: see ``_split``."""
    if dual:
        return bn(conv(x, bn_stats=True), relu=relu, residual=residual, dual=True)
    return bn(conv(x, bn_stats=True), relu=relu, residual=residual)


_DUAL = os.environ.get("PDL_RESNET_DUAL_OUTPUT", "1") != "0"


def _split(x):
    """A block's output feeds two consumers in the next block (its first convolution and its skip connection or downsample
    branch), so in backward two gradients come back for it.  Blocks therefore hand their output on as a PAIR of aliases of
    the same memory -- one per consumer -- and the closing BatchNorm's backward kernel adds the two gradients while it reads
    them (``pdl_bn_bwd2``) instead of autograd running an add kernel over two full tensors first (16 per ResNet-50 step)."""
    return x if isinstance(x, tuple) else (x, x)


def _open_block(conv, bn, x, skip=None):
    """First conv+bn+relu of an identity block, and the skip connection.  With PDL_SKIP_VIA_CONV=1 the skip is routed
    through the convolution's autograd node, so that in backward its gradient is added in that convolution's dgrad
    epilogue instead of by autograd's separate pass over two full tensors.  Measured on B200 (ResNet-50, batch 64):
    13.32 ms/step with the fused addend vs 13.08 ms with autograd's add -- the per-lane reads of the addend lengthen the
    epilogue of exactly the kernels that are epilogue-bound -- so it stays opt-in."""
    if os.environ.get("PDL_SKIP_VIA_CONV", "0") == "1":
        y, skip = conv(x, bn_stats=True, skip=True)
        return bn(y, relu=True), skip
    return bn(conv(x, bn_stats=True), relu=True), (x if skip is None else skip)


def _downsample(ds, x):
    """The ``downsample`` branch is ``Sequential(conv1x1, BatchNorm2d)`` (torchvision's layout, kept for the state_dict keys)."""
    if isinstance(ds, nn.Sequential) and len(ds) == 2 and isinstance(ds[0], nn.Conv2d) and isinstance(ds[1], nn.BatchNorm2d):
        return _conv_bn(ds[0], ds[1], x)
    return ds(x)


class BasicBlock(nn.Module):
    expansion = 1

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = _conv3x3(inplanes, planes, stride)
        self.bn1 = nn.BatchNorm2d(planes)
        self.relu = nn.ReLU(inplace=True)
        self.conv2 = _conv3x3(planes, planes)
        self.bn2 = nn.BatchNorm2d(planes)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        x, skip = _split(x)
        if self.downsample is None:
            out, identity = _open_block(self.conv1, self.bn1, x, skip)
        else:
            identity = _downsample(self.downsample, skip)
            out = _conv_bn(self.conv1, self.bn1, x, relu=True)
        return _conv_bn(self.conv2, self.bn2, out, relu=True, residual=identity, dual=_DUAL)


class Bottleneck(nn.Module):
    expansion = 4

    def __init__(self, inplanes, planes, stride=1, downsample=None):
        super().__init__()
        self.conv1 = _conv1x1(inplanes, planes)
        self.bn1 = nn.BatchNorm2d(planes)
        self.conv2 = _conv3x3(planes, planes, stride)
        self.bn2 = nn.BatchNorm2d(planes)
        self.conv3 = _conv1x1(planes, planes * self.expansion)
        self.bn3 = nn.BatchNorm2d(planes * self.expansion)
        self.relu = nn.ReLU(inplace=True)
        self.downsample = downsample
        self.stride = stride

    def forward(self, x):
        x, skip = _split(x)
        if self.downsample is None:
            out, identity = _open_block(self.conv1, self.bn1, x, skip)
        else:
            identity = _downsample(self.downsample, skip)
            out = _conv_bn(self.conv1, self.bn1, x, relu=True)
        out = _conv_bn(self.conv2, self.bn2, out, relu=True)
        return _conv_bn(self.conv3, self.bn3, out, relu=True, residual=identity, dual=_DUAL)


class ResNet(nn.Module):
    def __init__(self, block, layers, num_classes=1000):
        super().__init__()
        self.inplanes = 64
        self.conv1 = nn.Conv2d(3, 64, kernel_size=7, stride=2, padding=3, bias=False)
        self.bn1 = nn.BatchNorm2d(64)
        self.relu = nn.ReLU(inplace=True)
        self.maxpool = nn.MaxPool2d(kernel_size=3, stride=2, padding=1)
        self.layer1 = self._make_layer(block, 64, layers[0])
        self.layer2 = self._make_layer(block, 128, layers[1], stride=2)
        self.layer3 = self._make_layer(block, 256, layers[2], stride=2)
        self.layer4 = self._make_layer(block, 512, layers[3], stride=2)
        self.avgpool = nn.AdaptiveAvgPool2d((1, 1))
        self.fc = nn.Linear(512 * block.expansion, num_classes)
        for m in self.modules():
            if isinstance(m, nn.Conv2d):
                nn.init.kaiming_normal_(m.weight, mode="fan_out", nonlinearity="relu")
            elif isinstance(m, nn.BatchNorm2d):
                nn.init.constant_(m.weight, 1)
                nn.init.constant_(m.bias, 0)

    def _make_layer(self, block, planes, blocks, stride=1):
        downsample = None
        if stride != 1 or self.inplanes != planes * block.expansion:
            downsample = nn.Sequential(_conv1x1(self.inplanes, planes * block.expansion, stride), nn.BatchNorm2d(planes * block.expansion))
        layers = [block(self.inplanes, planes, stride, downsample)]
        self.inplanes = planes * block.expansion
        for _ in range(1, blocks):
            layers.append(block(self.inplanes, planes))
        return nn.Sequential(*layers)

    def forward(self, x):
        x = self.maxpool(_conv_bn(self.conv1, self.bn1, x, relu=True))
        x = self.layer4(self.layer3(self.layer2(self.layer1(x))))
        if isinstance(x, tuple):
            x = x[0]  # the last block's output has a single consumer
        x = self.avgpool(x)
        x = x.flatten(1)
        return self.fc(x)


def resnet18(pretrained=False, num_classes=1000, **kwargs):
    if pretrained:
        raise ValueError("no pretrained weights are shipped (and the training boxes have no network)")
    return ResNet(BasicBlock, [2, 2, 2, 2], num_classes=num_classes)


def resnet34(pretrained=False, num_classes=1000, **kwargs):
    if pretrained:
        raise ValueError("no pretrained weights are shipped (and the training boxes have no network)")
    return ResNet(BasicBlock, [3, 4, 6, 3], num_classes=num_classes)


def resnet50(pretrained=False, num_classes=1000, **kwargs):
    if pretrained:
        raise ValueError("no pretrained weights are shipped (and the training boxes have no network)")
    return ResNet(Bottleneck, [3, 4, 6, 3], num_classes=num_classes)
