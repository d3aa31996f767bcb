"""Data-parallel strategy: process groups, DDP wrapper + reducer, sampler, launchers."""
from . import distributed  # noqa: F401
from .ddp import DistributedDataParallel, compute_bucket_assignment_by_size  # noqa: F401
from .sampler import DistributedSampler  # noqa: F401
from .spawn import spawn  # noqa: F401
