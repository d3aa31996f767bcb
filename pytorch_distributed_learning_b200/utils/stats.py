"""Per-step statistics: device-timed step, exposed communication, bus bandwidth."""
from __future__ import annotations

import torch


def busbw_gbs(nbytes, seconds, world):
    """Allreduce bus bandwidth (NCCL convention): S/t * 2(W-1)/W, in GB/s."""
    if world <= 1 or seconds <= 0:
        return 0.0
    return nbytes / seconds * 2.0 * (world - 1) / world / 1e9


class CudaStepTimer:
    """CUDA-event timer on the current stream; ``ms()`` synchronises."""

    def __init__(self):
        self.start = torch.cuda.Event(enable_timing=True)
        self.stop = torch.cuda.Event(enable_timing=True)

    def __enter__(self):
        self.start.record()
        return self

    def __exit__(self, *exc):
        self.stop.record()

    def ms(self):
        self.stop.synchronize()
        return self.start.elapsed_time(self.stop)


def max_over_ranks(value, group=None):
    """Max of a python float over the group (device-timed numbers are reported as the max over ranks)."""
    from ..parallel import distributed as dist

    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return float(value)
    pg = dist._resolve(group)
    return max(pg.allgather_object(float(value)))
