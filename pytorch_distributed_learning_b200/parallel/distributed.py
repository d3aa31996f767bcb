"""Process groups and the module-level collective API.

Keeps the six-call surface the reference teaches (README.md:38-43): ``init_process_group``,
``new_group``, ``get_rank`` / ``get_world_size``, collectives, ``destroy_process_group``.

Backends
  * ``"gloo"`` / ``"cpu"``   -> :class:`CpuProcessGroup` (C++ TCP mesh, csrc/host/cpu_backend.cpp)
  * ``"nccl"`` / ``"b200"``  -> :class:`~.b200_backend.B200ProcessGroup`: hand-written sm_100a
    kernels that read/write peers' symmetric memory over NVLink.  The string ``"nccl"`` is
    accepted so that the reference scripts (launch_dist.py:49, mpspawn_dist.py:50,
    example_mp.py:38, example_launch.py:20) run with only their import changed -- no NCCL
    call is made on that path.
"""
from __future__ import annotations

import os
import pickle
import socket
import sys
import threading
from datetime import timedelta

import torch

from .. import _lib
from .rendezvous import rendezvous
from .store import PrefixStore, TCPStore

__all__ = [
    "ReduceOp", "ProcessGroup", "CpuProcessGroup", "init_process_group", "is_initialized", "is_available",
    "destroy_process_group", "new_group", "get_rank", "get_world_size", "get_backend", "barrier", "all_reduce",
    "broadcast", "all_gather", "all_gather_object", "broadcast_object_list", "send", "recv", "GroupMember",
    "get_process_group_ranks", "is_nccl_available", "is_gloo_available", "is_mpi_available", "reduce",
]


class ReduceOp:
    SUM, PRODUCT, MIN, MAX, AVG = 0, 1, 2, 3, 4


class GroupMember:
    WORLD = None
    NON_GROUP_MEMBER = object()


_DTYPE_CODE = {torch.float32: 0, torch.float64: 1, torch.int32: 2, torch.int64: 3, torch.uint8: 4, torch.bool: 4, torch.int8: 4}

DEFAULT_PG_TIMEOUT = timedelta(minutes=30)
DEFAULT_GPU_TIMEOUT = timedelta(minutes=10)


def is_available():
    return True


def is_gloo_available():
    return True


def is_nccl_available():
    """True when the first-party NVLink backend (selected by the name 'nccl'/'b200') can run."""
    return torch.cuda.is_available() and _lib.cuda_available_lib()


def is_mpi_available():
    return False


class ProcessGroup:
    """A set of ranks plus a data plane.  ``rank``/``size`` are group-local."""

    backend = "undefined"

    def __init__(self, store, group_rank, ranks, name):
        self.store = store
        self._rank = group_rank
        self.ranks = list(ranks)  # global ranks, position = group rank
        self.name = name

    def rank(self):
        return self._rank

    def size(self):
        return len(self.ranks)

    def global_rank(self, group_rank):
        return self.ranks[group_rank]

    def group_rank_of(self, global_rank):
        return self.ranks.index(global_rank)

    # data plane, overridden by backends -------------------------------------------
    def allreduce(self, tensor, op=ReduceOp.SUM):
        raise NotImplementedError

    def broadcast(self, tensor, src):
        raise NotImplementedError

    def allgather(self, tensor):
        raise NotImplementedError

    def barrier(self):
        raise NotImplementedError

    def send(self, tensor, dst):
        raise NotImplementedError

    def recv(self, tensor, src):
        raise NotImplementedError

    def destroy(self):
        pass

    # object helpers (control plane) -----------------------------------------------
    def allgather_object(self, obj):
        data = pickle.dumps(obj)
        n = torch.tensor([len(data)], dtype=torch.int64)
        sizes = [int(t.item()) for t in self._cpu_allgather(n)]
        buf = torch.zeros(max(sizes), dtype=torch.uint8)
        buf[: len(data)] = torch.frombuffer(bytearray(data), dtype=torch.uint8)
        parts = self._cpu_allgather(buf)
        return [pickle.loads(bytes(p[:s].numpy().tobytes())) for p, s in zip(parts, sizes)]

    def broadcast_object(self, obj, src):
        if self._rank == src:
            data = pickle.dumps(obj)
            n = torch.tensor([len(data)], dtype=torch.int64)
        else:
            n = torch.zeros(1, dtype=torch.int64)
        self._cpu_broadcast(n, src)
        buf = torch.zeros(int(n.item()), dtype=torch.uint8)
        if self._rank == src:
            buf.copy_(torch.frombuffer(bytearray(data), dtype=torch.uint8))
        self._cpu_broadcast(buf, src)
        return obj if self._rank == src else pickle.loads(bytes(buf.numpy().tobytes()))

    def _cpu_allgather(self, t):
        return self.allgather(t)

    def _cpu_broadcast(self, t, src):
        return self.broadcast(t, src)


def _local_addr_towards(host):
    """IP other ranks can use to reach this process (loopback when the master is local)."""
    if host in ("127.0.0.1", "localhost", "::1"):
        return "127.0.0.1"
    try:
        s = socket.socket(socket.AF_INET, socket.SOCK_DGRAM)
        s.connect((host, 9))
        ip = s.getsockname()[0]
        s.close()
        return ip
    except OSError:
        return "127.0.0.1"


class CpuProcessGroup(ProcessGroup):
    backend = "gloo"

    def __init__(self, store, group_rank, ranks, name, master_host, timeout_s):
        super().__init__(store, group_rank, ranks, name)
        L = _lib.host()
        prefix = (getattr(store, "full_prefix", "") + "/cpu").encode()
        self._comm = L.pdl_cpu_comm_create(
            store.client_handle, prefix, group_rank, len(ranks), _local_addr_towards(master_host).encode(), float(timeout_s)
        )
        if not self._comm:
            raise RuntimeError("CPU backend: could not connect the ranks of group %r" % name)
        self._lock = threading.Lock()
        self._timeout_s = float(timeout_s)

    def _check(self, rc, what):
        if rc == -2:  # the group's timeout (init_process_group(timeout=...)) bounds every collective, like NCCL's watchdog
            raise RuntimeError("CPU backend: %s timed out after %.0f s (a peer is alive but did not join the collective)" % (what, self._timeout_s))
        if rc != 0:
            raise RuntimeError("CPU backend: %s failed (a peer exited or the connection dropped)" % what)

    def _contig_cpu(self, tensor):
        if tensor.device.type != "cpu":
            raise RuntimeError("the CPU backend only moves CPU tensors (got %s)" % tensor.device)
        return tensor if tensor.is_contiguous() else tensor.contiguous()

    def allreduce(self, tensor, op=ReduceOp.SUM):
        t = self._contig_cpu(tensor)
        if t.dtype not in _DTYPE_CODE:
            raise TypeError("allreduce: unsupported dtype %s" % t.dtype)
        code = ReduceOp.SUM if op == ReduceOp.AVG else op
        with self._lock:
            self._check(_lib.host().pdl_cpu_allreduce(self._comm, t.data_ptr(), t.numel(), _DTYPE_CODE[t.dtype], code), "allreduce")
        if op == ReduceOp.AVG:
            t.div_(self.size())
        if t is not tensor:
            tensor.copy_(t)
        return tensor

    def broadcast(self, tensor, src):
        t = self._contig_cpu(tensor)
        with self._lock:
            self._check(_lib.host().pdl_cpu_broadcast(self._comm, t.data_ptr(), t.numel() * t.element_size(), int(src)), "broadcast")
        if t is not tensor:
            tensor.copy_(t)
        return tensor

    def allgather(self, tensor):
        t = self._contig_cpu(tensor)
        out = torch.empty((self.size(),) + tuple(t.shape), dtype=t.dtype)
        with self._lock:
            self._check(_lib.host().pdl_cpu_allgather(self._comm, t.data_ptr(), out.data_ptr(), t.numel() * t.element_size()), "allgather")
        return list(out.unbind(0))

    def barrier(self):
        with self._lock:
            self._check(_lib.host().pdl_cpu_barrier(self._comm), "barrier")

    def send(self, tensor, dst):
        t = self._contig_cpu(tensor)
        with self._lock:
            self._check(_lib.host().pdl_cpu_send(self._comm, t.data_ptr(), t.numel() * t.element_size(), int(dst)), "send")

    def recv(self, tensor, src):
        t = self._contig_cpu(tensor)
        with self._lock:
            self._check(_lib.host().pdl_cpu_recv(self._comm, t.data_ptr(), t.numel() * t.element_size(), int(src)), "recv")
        if t is not tensor:
            tensor.copy_(t)
        return tensor

    def destroy(self):
        if self._comm:
            _lib.host().pdl_cpu_comm_destroy(self._comm)
            self._comm = None


# ---------------------------------------------------------------------------------------
# world state
# ---------------------------------------------------------------------------------------
class _World:
    def __init__(self):
        self.reset()

    def reset(self):
        self.default_pg = None
        self.store = None
        self.rank = -1
        self.world_size = -1
        self.backend = None
        self.master_host = "127.0.0.1"
        self.timeout_s = 1800.0
        self.group_count = 0
        self.groups = []
        self.prev_excepthook = None


_world = _World()


def _canonical_backend(backend):
    if backend is None:
        return "b200" if torch.cuda.is_available() else "gloo"
    b = str(backend).lower()
    if b in ("gloo", "cpu"):
        return "gloo"
    if b in ("nccl", "b200", "blackwell", "nvlink"):
        return "b200"
    if b == "mpi":
        raise RuntimeError("the MPI backend is not available in this build (is_mpi_available() is False)")
    raise ValueError("unknown backend %r (expected 'nccl'/'b200' or 'gloo'/'cpu')" % backend)


def _make_pg(backend, store, group_rank, ranks, name, device_index=None):
    if backend == "gloo":
        return CpuProcessGroup(store, group_rank, ranks, name, _world.master_host, _world.timeout_s)
    from .b200_backend import B200ProcessGroup

    return B200ProcessGroup(store, group_rank, ranks, name, _world.master_host, _world.timeout_s, device_index)


def is_initialized():
    return _world.default_pg is not None


def _install_excepthook(rank):
    prev = sys.excepthook

    def hook(tp, val, tb):
        import io
        import traceback

        buf = io.StringIO()
        traceback.print_exception(tp, val, tb, file=buf)
        prefix = "[rank%d]:" % rank
        sys.stderr.write("\n".join(prefix + line for line in buf.getvalue().split("\n")) + "\n")

    sys.excepthook = hook
    _world.prev_excepthook = prev


def init_process_group(backend=None, init_method=None, timeout=None, world_size=-1, rank=-1, store=None, group_name="", device_id=None):
    """Create the default (world) group.  Signature follows the reference's call sites:
    ``init_process_group(backend='nccl', init_method='env://'|'tcp://h:p', world_size=W, rank=r)``."""
    if is_initialized():
        raise ValueError("trying to initialize the default process group twice!")
    if store is not None and init_method is not None:
        raise ValueError("Cannot specify both init_method and store.")
    be = _canonical_backend(backend)
    if timeout is None:
        timeout = DEFAULT_GPU_TIMEOUT if be == "b200" else DEFAULT_PG_TIMEOUT
    timeout_s = timeout.total_seconds() if isinstance(timeout, timedelta) else float(timeout)
    if store is None:
        store, rank, world_size = rendezvous(init_method or "env://", rank, world_size, timeout=timeout_s)
        master_host = store.host
    else:
        if rank < 0 or world_size < 1:
            raise ValueError("rank and world_size are required when a store is passed")
        master_host = getattr(store, "host", "127.0.0.1")
    _world.store, _world.rank, _world.world_size = store, rank, world_size
    _world.backend, _world.master_host, _world.timeout_s = be, master_host, timeout_s
    dev_index = None
    if device_id is not None:
        dev_index = device_id.index if isinstance(device_id, torch.device) else int(device_id)
    try:
        pg = _make_pg(be, PrefixStore("default_pg", store), rank, list(range(world_size)), "default_pg", dev_index)
    except Exception:
        _world.reset()
        raise
    _world.default_pg = pg
    _world.groups = [pg]
    GroupMember.WORLD = pg
    _install_excepthook(rank)
    return pg


def _default():
    if _world.default_pg is None:
        raise ValueError("Default process group has not been initialized, please make sure to call init_process_group.")
    return _world.default_pg


def _resolve(group):
    return _default() if group is None else group


def get_rank(group=None):
    if group is GroupMember.NON_GROUP_MEMBER:
        return -1
    if group is None:
        _default()
        return _world.rank
    return group.rank()


def get_world_size(group=None):
    if group is GroupMember.NON_GROUP_MEMBER:
        return -1
    return _resolve(group).size()


def get_backend(group=None):
    pg = _resolve(group)
    return "nccl" if pg.backend == "b200" else pg.backend


def get_process_group_ranks(group=None):
    return list(_resolve(group).ranks)


def new_group(ranks=None, timeout=None, backend=None, **_unused):
    """Create a sub-group.  Like the reference's stack requires, EVERY rank of the world must
    call this (same order, same arguments), members and non-members alike."""
    _default()
    _world.group_count += 1
    gid = _world.group_count
    ranks = sorted(set(range(_world.world_size) if ranks is None else ranks))
    for r in ranks:
        if not (0 <= r < _world.world_size):
            raise ValueError("new_group: rank %d is not part of the world (size %d)" % (r, _world.world_size))
    if _world.rank not in ranks:
        return GroupMember.NON_GROUP_MEMBER
    be = _canonical_backend(backend) if backend is not None else _world.backend
    pg = _make_pg(be, PrefixStore("group_%d" % gid, _world.store), ranks.index(_world.rank), ranks, "group_%d" % gid)
    _world.groups.append(pg)
    return pg


def destroy_process_group(group=None):
    if group is GroupMember.NON_GROUP_MEMBER:
        return
    if group is None or group is _world.default_pg:
        for pg in reversed(_world.groups):
            pg.destroy()
        store = _world.store
        if _world.prev_excepthook is not None:
            sys.excepthook = _world.prev_excepthook
        rank, ws = _world.rank, _world.world_size
        _world.reset()
        GroupMember.WORLD = None
        if isinstance(store, TCPStore):
            # the server must outlive every client: non-masters check out first
            try:
                if not store.is_master:
                    store.add("__shutdown__", 1)
                else:
                    import time

                    deadline = time.time() + 10.0
                    while ws > 1 and store.add("__shutdown__", 0) < ws - 1 and time.time() < deadline:
                        time.sleep(0.01)
            except Exception:
                pass
            store.close()
    else:
        group.destroy()
        if group in _world.groups:
            _world.groups.remove(group)


# -- collectives --------------------------------------------------------------------------
def _skip(group):
    return group is GroupMember.NON_GROUP_MEMBER


def all_reduce(tensor, op=ReduceOp.SUM, group=None, async_op=False):
    if _skip(group):
        return None
    _resolve(group).allreduce(tensor, op)
    return _DoneWork() if async_op else None


def broadcast(tensor, src, group=None, async_op=False):
    """``src`` is a GLOBAL rank, as in torch."""
    if _skip(group):
        return None
    pg = _resolve(group)
    pg.broadcast(tensor, pg.group_rank_of(src))
    return _DoneWork() if async_op else None


def reduce(tensor, dst, op=ReduceOp.SUM, group=None):
    """Result is only guaranteed on ``dst`` (implemented as an allreduce)."""
    if _skip(group):
        return None
    _resolve(group).allreduce(tensor, op)


def all_gather(tensor_list, tensor, group=None):
    if _skip(group):
        return None
    parts = _resolve(group).allgather(tensor)
    for dst, src in zip(tensor_list, parts):
        dst.copy_(src)


def all_gather_into_tensor(output_tensor, input_tensor, group=None):
    """``output_tensor`` (world x the input's size along dim 0, or stacked) receives every rank's input, in rank order."""
    if _skip(group):
        return None
    parts = _resolve(group).allgather(input_tensor)
    flat = output_tensor.view(len(parts), -1)
    for i, src in enumerate(parts):
        flat[i].copy_(src.reshape(-1))


def reduce_scatter_tensor(output, input, op=ReduceOp.SUM, group=None):
    """Rank r receives the r-th of ``world`` equal chunks of the element-wise reduction of ``input`` over the group
    (composed from the group's allreduce: every variant of it is bit-identical across ranks, so this is too)."""
    if _skip(group):
        return None
    pg = _resolve(group)
    world = len(pg.ranks)
    if input.numel() != output.numel() * world:
        raise ValueError("reduce_scatter_tensor: input must hold world_size (%d) chunks of the output's size" % world)
    full = input.detach().clone().contiguous()
    pg.allreduce(full, op)
    output.copy_(full.view(world, -1)[pg.rank()].view_as(output))


def gather(tensor, gather_list=None, dst=0, group=None):
    """``gather_list`` (on global rank ``dst`` only) receives every rank's tensor."""
    if _skip(group):
        return None
    pg = _resolve(group)
    parts = pg.allgather(tensor)
    if pg.rank() == pg.group_rank_of(dst):
        if gather_list is None or len(gather_list) != len(parts):
            raise ValueError("gather: the destination rank must pass a gather_list of world_size tensors")
        for d, s in zip(gather_list, parts):
            d.copy_(s)


def scatter(tensor, scatter_list=None, src=0, group=None):
    """Rank r receives ``scatter_list[r]`` of global rank ``src``."""
    if _skip(group):
        return None
    pg = _resolve(group)
    world = len(pg.ranks)
    stacked = torch.empty((world,) + tuple(tensor.shape), dtype=tensor.dtype, device=tensor.device)
    if pg.rank() == pg.group_rank_of(src):
        if scatter_list is None or len(scatter_list) != world:
            raise ValueError("scatter: the source rank must pass a scatter_list of world_size tensors")
        for i, s in enumerate(scatter_list):
            stacked[i].copy_(s)
    pg.broadcast(stacked, pg.group_rank_of(src))
    tensor.copy_(stacked[pg.rank()])


def all_gather_object(object_list, obj, group=None):
    if _skip(group):
        return None
    for i, o in enumerate(_resolve(group).allgather_object(obj)):
        object_list[i] = o


def broadcast_object_list(object_list, src=0, group=None):
    if _skip(group):
        return None
    pg = _resolve(group)
    out = pg.broadcast_object(list(object_list), pg.group_rank_of(src))
    object_list[:] = out


def barrier(group=None, device_ids=None):
    if _skip(group):
        return None
    _resolve(group).barrier()


def send(tensor, dst, group=None):
    pg = _resolve(group)
    pg.send(tensor, pg.group_rank_of(dst))


def recv(tensor, src, group=None):
    pg = _resolve(group)
    pg.recv(tensor, pg.group_rank_of(src))


class _DoneWork:
    def wait(self, timeout=None):
        return True

    def is_completed(self):
        return True
