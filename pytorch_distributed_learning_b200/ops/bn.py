"""BatchNorm2d (training-mode batch statistics + running-stat update), NHWC."""
from __future__ import annotations

import torch.nn.functional as F


def batch_norm(x, weight, bias, running_mean, running_var, training, momentum, eps):
    return F.batch_norm(x, running_mean, running_var, weight, bias, training, momentum, eps)
