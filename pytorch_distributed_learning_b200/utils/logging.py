"""Rank-gated printing, as the reference does by hand (``if local_rank == 0: print`` in
launch_dist.py:93-101, ``if rank == 0`` in example_launch.py:95), plus PDL_DEBUG tracing --
the analogue of the ``NCCL_DEBUG=INFO`` tip in the reference README (README.md:218-221)."""
from __future__ import annotations

import os
import sys
import time

from ..parallel import distributed as dist


def debug_level():
    v = os.environ.get("PDL_DEBUG", "").upper()
    return {"": 0, "0": 0, "WARN": 1, "1": 1, "INFO": 2, "2": 2, "TRACE": 3, "3": 3}.get(v, 0)


def debug(msg, level=2):
    if debug_level() >= level:
        r = dist.get_rank() if dist.is_initialized() else -1
        sys.stderr.write("[pdl rank%d %.3f] %s\n" % (r, time.time() % 1000, msg))
        sys.stderr.flush()


def print_rank0(*args, local=False, **kwargs):
    """print on global rank 0 (or on every node's local rank 0 with ``local=True``)."""
    if local:
        ok = int(os.environ.get("LOCAL_RANK", "0")) == 0
    else:
        ok = (not dist.is_initialized()) or dist.get_rank() == 0
    if ok:
        print(*args, **kwargs)


class StepTimer:
    """Wall-clock bracket like the reference's ``datetime.now()`` delta (launch_dist.py:76,102)."""

    def __init__(self):
        self.t0 = time.time()

    def elapsed(self):
        return time.time() - self.t0
