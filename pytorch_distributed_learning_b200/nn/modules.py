"""Layer modules backed by this package's sm_100a kernels.

Each class subclasses its ``torch.nn`` namesake (same constructor, same parameters, same
``state_dict`` keys) and overrides ``forward``: fp32 CUDA tensors go to the native kernels in
``ops.functional`` (channels-last/NHWC internally); anything else (CPU tensors in the unit
tests, unsupported dtypes) uses the stock implementation.  Layers covered are the ones the
reference's models use: Conv2d/ReLU/MaxPool2d/Linear/CrossEntropyLoss (launch_dist.py:9-41,58)
and BatchNorm2d/AdaptiveAvgPool2d for the torchvision ResNets (example_launch.py:26).
"""
from __future__ import annotations

import torch
import torch.nn as tnn


def _native(x):
    return x.is_cuda and x.dtype == torch.float32


def _pair(v):
    return tuple(v) if isinstance(v, (tuple, list)) else (v, v)


class Conv2d(tnn.Conv2d):
    """``conv(x)``.  ``conv(x, bn_stats=True)`` tells the tensor-core kernel that a training-mode ``BatchNorm2d`` consumes
    the result: its epilogue then also reduces the per-channel statistics of the output tile and the result carries them
    (``_pdl_bn_partial``), so the normalisation layer does not have to re-read the tensor.  A hint, never a requirement.
    ``conv(x, skip=True)`` returns ``(y, x_skip)``: ``x_skip`` is x, routed through this layer's autograd node, for use as
    the residual of the block this convolution opens -- its gradient is then added inside the dgrad kernel."""

    def forward(self, x, bn_stats=False, skip=False):
        if (_native(x) and self.groups == 1 and _pair(self.dilation) == (1, 1) and self.padding_mode == "zeros"
                and not isinstance(self.padding, str) and self.stride[0] == self.stride[1] and self.padding[0] == self.padding[1]
                and self.kernel_size[0] == self.kernel_size[1]):
            from ..ops import functional as NF

            want_stats = bn_stats and self.bias is None and torch.is_grad_enabled()
            if skip and torch.is_grad_enabled() and x.requires_grad:
                y, partial, x_skip = NF.conv2d(x, self.weight, self.bias, self.stride[0], self.padding[0], bn_stats=want_stats, with_skip=True)
                if partial is not None:
                    y._pdl_bn_partial = partial
                return y, x_skip
            if want_stats:
                y, partial = NF.conv2d(x, self.weight, None, self.stride[0], self.padding[0], bn_stats=True)
                if partial is not None:
                    y._pdl_bn_partial = partial
                return (y, x) if skip else y
            y = NF.conv2d(x, self.weight, self.bias, self.stride[0], self.padding[0])
            return (y, x) if skip else y
        y = super().forward(x)
        return (y, x) if skip else y


class ReLU(tnn.ReLU):
    def forward(self, x):
        if _native(x):
            from ..ops import functional as NF

            return NF.relu(x)
        return super().forward(x)


class MaxPool2d(tnn.MaxPool2d):
    def forward(self, x):
        k, s, p = _pair(self.kernel_size), _pair(self.stride if self.stride is not None else self.kernel_size), _pair(self.padding)
        if (_native(x) and k[0] == k[1] and s[0] == s[1] and p[0] == p[1] and _pair(self.dilation) == (1, 1)
                and not self.ceil_mode and not self.return_indices):
            from ..ops import functional as NF

            return NF.max_pool2d(x, k[0], s[0], p[0])
        return super().forward(x)


class Linear(tnn.Linear):
    def forward(self, x):
        if _native(x) and x.dim() == 2:
            from ..ops import functional as NF

            return NF.linear(x, self.weight, self.bias)
        return super().forward(x)


class CrossEntropyLoss(tnn.CrossEntropyLoss):
    def forward(self, logits, target):
        from ..parallel.lazy_step import DeferredLoss, DeferredOutput

        if isinstance(logits, DeferredOutput):
            # the model's forward pass has not run yet (parallel/lazy_step.py): keep deferring when this is the loss the graphed
            # step computes, otherwise ask for the real logits now
            if (logits._value is None and isinstance(target, torch.Tensor) and target.dtype == torch.int64 and target.dim() == 1
                    and self.weight is None and self.reduction == "mean" and self.label_smoothing == 0.0 and self.ignore_index == -100):
                return DeferredLoss(logits, target, self)
            logits = logits._tensor()
        if (_native(logits) and logits.dim() == 2 and target.dtype == torch.int64 and self.weight is None
                and self.reduction == "mean" and self.label_smoothing == 0.0 and self.ignore_index == -100):
            from ..ops import functional as NF

            return NF.cross_entropy(logits, target)
        return super().forward(logits, target)


class BatchNorm2d(tnn.BatchNorm2d):
    """``bn(x)``; the optional arguments fuse what follows it in a ResNet block into the same kernels:
    ``bn(x, relu=True)`` = ``relu(bn(x))`` and ``bn(x, relu=True, residual=r)`` = ``relu(bn(x) + r)``."""

    def forward(self, x, relu=False, residual=None, dual=False):
        """``dual=True`` returns ``(y, y')``: the output as two tensor objects over the same memory, for a consumer pair such
        as the next residual block's first convolution and its skip connection (their two gradients are then added inside
        this layer's backward kernel instead of by a separate autograd pass)."""
        if _native(x) and self.affine and self.track_running_stats and self.momentum is not None and (residual is None or _native(residual)):
            from ..ops import functional as NF

            return NF.batch_norm(x, self.weight, self.bias, self.running_mean, self.running_var, self.training, self.momentum, self.eps,
                                 relu, residual, self.num_batches_tracked,  # the step counter is bumped by the finalize kernel
                                 partial=getattr(x, "_pdl_bn_partial", None) if self.training else None, dual=dual)
        y = super().forward(x)
        if residual is not None:
            y = y + residual
        y = torch.relu(y) if relu else y
        return (y, y) if dual else y


class AdaptiveAvgPool2d(tnn.AdaptiveAvgPool2d):
    def forward(self, x):
        if _native(x) and _pair(self.output_size) == (1, 1):
            from ..ops import functional as NF

            return NF.global_avg_pool(x)
        return super().forward(x)


def add_relu(a, b):
    """relu(a + b), the residual join of a ResNet block: one native kernel on fp32 CUDA tensors."""
    if _native(a) and _native(b) and a.shape == b.shape:
        from ..ops import functional as NF

        return NF.add_relu(a, b)
    return torch.relu(a + b)
