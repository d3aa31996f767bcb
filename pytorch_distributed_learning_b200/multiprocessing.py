"""``pdl.multiprocessing.spawn`` -- same call shape as the ``mp.spawn`` the reference uses
(mpspawn_dist.py:140, example_mp.py:27)."""
from .parallel.spawn import (ProcessContext, ProcessExitedException, ProcessRaisedException,  # noqa: F401
                             spawn, start_processes)
