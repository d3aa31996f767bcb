from torch.utils.data import DataLoader, Dataset  # noqa: F401

from . import datasets, distributed, transforms  # noqa: F401
from .datasets import CIFAR10, MNIST  # noqa: F401
from .prefetch import CudaPrefetcher  # noqa: F401
from .synthetic import SyntheticCIFAR10, SyntheticDataset, SyntheticImageNet, SyntheticMNIST  # noqa: F401
