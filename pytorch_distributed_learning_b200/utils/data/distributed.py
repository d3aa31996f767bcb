"""``utils.data.distributed.DistributedSampler`` (reference launch_dist.py:67 spelling)."""
from ...parallel.sampler import DistributedSampler  # noqa: F401
