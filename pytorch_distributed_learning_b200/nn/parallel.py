"""``nn.parallel.DistributedDataParallel`` (reference launch_dist.py:61 spelling)."""
from ..parallel.ddp import DistributedDataParallel  # noqa: F401
