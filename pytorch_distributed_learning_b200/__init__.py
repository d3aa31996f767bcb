"""B200-native data-parallel training framework.

User surface (what the reference's four scripts touch, one import away):

    import pytorch_distributed_learning_b200 as pdl
    pdl.distributed.init_process_group(backend="nccl", init_method="env://" | "tcp://host:port", world_size=W, rank=r)
    pdl.distributed.new_group(ranks) / get_rank() / get_world_size() / destroy_process_group()
    pdl.nn.parallel.DistributedDataParallel(model, device_ids=[g], output_device=g)
    pdl.utils.data.distributed.DistributedSampler(dataset, num_replicas=W, rank=r)
    pdl.multiprocessing.spawn(fn, nprocs=G, args=(...))          # launch path 1
    python -m pytorch_distributed_learning_b200.launch ...      # launch path 2
    pdl.optim.SGD(...)                                           # update fused into the allreduce kernel
    pdl.TrainStep(ddp_model, criterion, optimizer)               # whole step as one CUDA graph
    pdl.utils.checkpoint.save_checkpoint / load_checkpoint       # rank-0 state_dict files
"""
from __future__ import annotations

__version__ = "0.1.0"

from . import _build, _lib  # noqa: F401
from .parallel import distributed  # noqa: F401
from .parallel.ddp import DistributedDataParallel  # noqa: F401
from .parallel.sampler import DistributedSampler  # noqa: F401
from .parallel.spawn import spawn  # noqa: F401
from . import multiprocessing  # noqa: F401
from . import optim  # noqa: F401
from . import nn  # noqa: F401
from . import utils  # noqa: F401
from . import models  # noqa: F401
from .parallel.train_step import TrainStep  # noqa: F401
from .parallel.pipeline import PendingLoss  # noqa: F401

init_process_group = distributed.init_process_group
destroy_process_group = distributed.destroy_process_group
new_group = distributed.new_group
get_rank = distributed.get_rank
get_world_size = distributed.get_world_size
