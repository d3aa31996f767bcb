"""Drop-in ``nn`` namespace: swap ``import torch.nn as nn`` for
``import pytorch_distributed_learning_b200.nn as nn`` and the reference's ``ConvNet``
(launch_dist.py:9-41) is built from this package's sm_100a kernels instead of cuDNN/cuBLAS.

Everything not overridden here is re-exported from ``torch.nn`` unchanged.
"""
from torch.nn import *  # noqa: F401,F403
from torch.nn import Module, Parameter, Sequential, init  # noqa: F401

import importlib as _importlib

# `from torch.nn import *` already bound the name `parallel` (torch's); bind ours explicitly
parallel = _importlib.import_module(__name__ + ".parallel")
from .modules import (AdaptiveAvgPool2d, BatchNorm2d, Conv2d, CrossEntropyLoss, Linear, MaxPool2d, ReLU)  # noqa: F401
