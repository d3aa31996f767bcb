from torch.utils.data import DataLoader, Dataset  # noqa: F401

from . import distributed  # noqa: F401
from .synthetic import SyntheticCIFAR10, SyntheticDataset, SyntheticImageNet, SyntheticMNIST  # noqa: F401
