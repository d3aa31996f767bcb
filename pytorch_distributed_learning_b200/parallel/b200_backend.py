"""The NVLink data plane: symmetric-memory segments + hand-written peer-to-peer collectives.

``B200ProcessGroup`` is what ``init_process_group(backend="nccl")`` creates in this package
(reference call sites launch_dist.py:49, mpspawn_dist.py:49-54, example_mp.py:37-42).  It makes
no NCCL call: buckets live in symmetric memory (csrc/cuda/symm.cu) and every collective is one
of the kernels in csrc/cuda/allreduce.cu.  A small TCP mesh (the CPU backend) rides along as the
control plane for object exchange and host-side barriers.

CUDA state is created lazily, on the first GPU collective or DDP wrap: the reference's spawn
scripts call ``torch.cuda.set_device`` only AFTER ``init_process_group`` (mpspawn_dist.py:49-60).
"""
from __future__ import annotations

import ctypes
import weakref
import os
import uuid

import torch

from .. import _lib
from . import distributed as dist
from ..utils.logging import debug

ALGOS = {"oneshot": 0, "twoshot": 1, "ring": 2, "nvls": 3, "push": 4}
ALGO_NAMES = {v: k for k, v in ALGOS.items()}


def _check(rc, what):
    if rc != 0:
        raise RuntimeError("%s failed (%d): %s" % (what, rc, _lib.cuda().pdl_last_error().decode()))


class _RawCuda:
    """Exposes a raw device range to torch (zero-copy) through __cuda_array_interface__."""

    def __init__(self, ptr, nbytes, owner):
        self.__cuda_array_interface__ = {"shape": (int(nbytes),), "typestr": "|u1", "data": (int(ptr), False), "version": 3, "strides": None}
        self._owner = owner


class SymmetricSegment:
    """``nbytes`` of device memory at the same offset on every rank of ``pg``, peer-mapped everywhere.

    Collective constructor.  Layout is the caller's business; ``tensor()`` returns torch views of
    the local copy, ``ptr(rank)`` the address through which THIS process reaches rank's copy."""

    def __init__(self, pg, nbytes, device_index, want_multicast=True):
        L = _lib.cuda()
        self.pg, self.device_index = pg, device_index
        self.world, self.rank = pg.size(), pg.rank()
        token = pg._ctl.broadcast_object(uuid.uuid4().hex if self.rank == 0 else None, 0)
        want_mc = bool(want_multicast and self.world > 1 and os.environ.get("PDL_DISABLE_NVLS", "0") != "1"
                       and L.pdl_symm_multicast_supported(device_index))
        want_mc = all(pg._ctl.allgather_object(want_mc))
        h = ctypes.c_void_p()
        _check(L.pdl_symm_begin(self.rank, self.world, device_index, int(nbytes), ("pdl-symm-" + token).encode(), int(want_mc), ctypes.byref(h)),
               "symmetric segment allocation")
        self._h = h
        pg._ctl.barrier()  # every rank is listening
        _check(L.pdl_symm_exchange(h), "symmetric handle exchange")
        mc_ok = all(pg._ctl.allgather_object(bool(L.pdl_symm_multicast_ready(h))))
        L.pdl_symm_bind_multicast(h, int(mc_ok))
        mc_ok = all(pg._ctl.allgather_object(bool(L.pdl_symm_mc_ptr(h))))
        if not mc_ok:
            L.pdl_symm_bind_multicast(h, 0)
        self.bytes = int(L.pdl_symm_bytes(h))
        self.ptrs = [int(L.pdl_symm_ptr(h, r) or 0) for r in range(self.world)]
        self.mc_ptr = int(L.pdl_symm_mc_ptr(h) or 0) if mc_ok else 0
        self._base = torch.as_tensor(_RawCuda(self.ptrs[self.rank], self.bytes, self), device="cuda:%d" % device_index)
        pg._ctl.barrier()  # everyone mapped (memory is zero-filled by begin)
        debug("symmetric segment: %d bytes, multicast=%s" % (self.bytes, bool(self.mc_ptr)))

    def tensor(self, offset_bytes, numel, dtype=torch.float32):
        nbytes = numel * torch.empty((), dtype=dtype).element_size()
        return self._base[offset_bytes:offset_bytes + nbytes].view(dtype)

    def ptr(self, rank, offset_bytes=0):
        return self.ptrs[rank] + offset_bytes

    def destroy(self):
        if self._h:
            self._base = None
            _lib.cuda().pdl_symm_destroy(self._h)
            self._h = None


def _ptr_array(values):
    return (ctypes.c_void_p * len(values))(*values)


_TABLE = None


def _load_table():
    """The measured selection table (``allreduce_table.json`` next to this file; written by ``bench/allreduce_sweep.py
    --emit-table`` on an 8 x B200 NVSwitch box): per world size, ascending size classes, each with the fastest variant and
    its CTA count -- with NVSwitch multicast and without.  ``PDL_ALLREDUCE_TABLE`` points at another file."""
    global _TABLE
    if _TABLE is None:
        import json

        path = os.environ.get("PDL_ALLREDUCE_TABLE") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "allreduce_table.json")
        with open(path) as f:
            raw = json.load(f)
        _TABLE = {int(w): rows for w, rows in raw["worlds"].items()}
    return _TABLE


def table_row(nbytes, world):
    table = _load_table()
    worlds = sorted(table)
    w = next((x for x in worlds if x >= world), worlds[-1])  # a world size that was not measured uses the next larger one
    for row in table[w]:
        if row["max_bytes"] is None or nbytes <= row["max_bytes"]:
            return row
    return table[w][-1]


def choose_algorithm(nbytes, world, has_multicast, override=None, table=True):
    """Per-bucket variant from MEASURED bus bandwidth (north star; SURVEY 5.9 "Selection"): the sweep's table says, per world
    size and size class, which kernel was fastest on this fabric.  ``override`` (``PDL_ALGO`` /
    ``DistributedDataParallel(algorithm=...)``) wins; ``table=False`` is the rule of thumb used for the process group's own
    staged collectives (control-plane traffic, not tuned)."""
    if override:
        name = override.lower()
        if name not in ALGOS:
            raise ValueError("unknown allreduce algorithm %r (choose from %s)" % (override, sorted(ALGOS)))
        if name == "nvls" and not has_multicast:
            raise RuntimeError("allreduce algorithm 'nvls' was requested but this box exposes no NVSwitch multicast")
        return ALGOS[name]
    if world == 1:
        return ALGOS["oneshot"]
    if table and os.environ.get("PDL_ALLREDUCE_TABLE", "") != "off":
        row = table_row(nbytes, world)
        return ALGOS[row["algo"] if has_multicast else row["algo_no_mc"]]
    if has_multicast:
        return ALGOS["nvls"]
    threshold = int(os.environ.get("PDL_ONESHOT_MAX_BYTES", 32 << 10))
    return ALGOS["oneshot"] if nbytes <= threshold else ALGOS["twoshot"]


def choose_blocks(nbytes, world, algo, has_multicast, override=None):
    """CTA count for a bucket kernel: the table's tuned value when the table chose the variant, else the library's rule."""
    n4 = (nbytes + 15) // 16
    L = _lib.cuda()
    if not override and world > 1 and os.environ.get("PDL_ALLREDUCE_TABLE", "") != "off":
        row = table_row(nbytes, world)
        name, blocks = (row["algo"], row.get("blocks")) if has_multicast else (row["algo_no_mc"], row.get("blocks_no_mc"))
        if ALGOS[name] == algo and blocks:
            return max(1, min(int(blocks), L.pdl_comm_max_blocks(), int(n4)))
    return L.pdl_allreduce_blocks(n4, world, algo)


class B200ProcessGroup(dist.ProcessGroup):
    backend = "b200"
    STAGING_BYTES = 64 << 20
    PUSH_STAGE_BYTES = 256 << 10  # per-pass payload of the push variant in pg.allreduce (its inbox holds 2 x W rows of this size)

    def __init__(self, store, group_rank, ranks, name, master_host, timeout_s, device_index=None):
        super().__init__(store, group_rank, ranks, name)
        if not torch.cuda.is_available():
            raise RuntimeError("backend 'nccl'/'b200' needs a CUDA device; use backend='gloo' for CPU-only runs")
        _lib.cuda()  # fail loudly now if the kernel library was not built
        self._ctl = dist.CpuProcessGroup(store, group_rank, ranks, name + "/ctl", master_host, timeout_s)
        # device-side bound on every peer wait (a dead rank must not spin a kernel for the 10-minute host timeout)
        self.timeout_s = min(float(timeout_s), float(os.environ.get("PDL_KERNEL_TIMEOUT_S", "60")))
        self.device_index = device_index
        # staging window of the group's own collectives (broadcast / allgather / allreduce of arbitrary tensors): 2 halves.  The
        # world group gets 64 MiB halves; a sub-group (new_group) 8 MiB, so that every extra group does not pin 128 MiB.
        default_mb = 64 if name in ("default_pg", "", None) or "default" in str(name) else 8
        self.STAGING_BYTES = int(os.environ.get("PDL_STAGING_MB", default_mb)) << 20
        self._staging = None
        self._segments = []
        self._comm_stream = None
        self._comm_states = []
        self._host_errors = None
        self._host_error_slots = 0
        if device_index is not None:
            self._ensure_cuda(device_index)

    # -- lazy CUDA bring-up -------------------------------------------------------------
    def _ensure_cuda(self, device_index=None):
        if self.device_index is None:
            self.device_index = torch.cuda.current_device() if device_index is None else device_index
        elif device_index is not None and device_index != self.device_index:
            raise RuntimeError("process group is bound to cuda:%d but was used with cuda:%d" % (self.device_index, device_index))
        if self._comm_stream is None:
            with torch.cuda.device(self.device_index):
                self._comm_stream = torch.cuda.Stream(priority=-1)
        return self.device_index

    @property
    def comm_stream(self):
        self._ensure_cuda()
        return self._comm_stream

    def new_segment(self, nbytes, want_multicast=True):
        seg = SymmetricSegment(self, nbytes, self._ensure_cuda(), want_multicast)
        self._segments.append(seg)
        return seg

    def new_comm_state(self):
        """Device-side state of one communicator (epochs, retire counters, the sticky error word).  The error word is
        mirrored into a pinned host word that the failing kernel writes itself, so ``poll_errors()`` costs a host load."""
        L = _lib.cuda()
        st = torch.zeros(L.pdl_comm_state_bytes(), dtype=torch.uint8, device="cuda:%d" % self._ensure_cuda())
        if self._host_errors is None:
            self._host_errors = torch.zeros(64, dtype=torch.int32).pin_memory()
            self._host_error_slots = 0
        if self._host_error_slots < self._host_errors.numel():
            slot = self._host_error_slots
            self._host_error_slots += 1
            off = L.pdl_comm_state_host_error_offset()
            st[off:off + 8].view(torch.int64).fill_(self._host_errors.data_ptr() + 4 * slot)
        self._comm_states.append(weakref.ref(st))  # every communicator state of this group: checked by check_device_errors()
        return st

    def poll_errors(self):
        """Cheap failure check for the step path (no synchronisation, no copy): raises once a kernel of this group has
        reported a lost peer through its pinned host word.  Kernels launched after the failure leave gradients and
        parameters untouched, so the replica stays at its last consistent state."""
        if self._host_errors is not None and bool(self._host_errors.any()):
            raise RuntimeError("[rank %d] a device collective of group %r timed out waiting for a peer GPU (PDL_KERNEL_TIMEOUT_S=%gs): a rank "
                               "died or hung.  Parameters were left as they were before the failed step." % (self.rank(), self.name, self.timeout_s))

    def _stage(self):
        if self._staging is None:
            L = _lib.cuda()
            pad = (L.pdl_signal_pad_bytes() + 4095) // 4096 * 4096
            half = self.STAGING_BYTES
            seg = self.new_segment(pad + 2 * half, want_multicast=False)
            self._staging = {"seg": seg, "pad": pad, "half": half, "state": self.new_comm_state(),
                             "pads": _ptr_array([seg.ptr(r) for r in range(self.size())])}
        return self._staging

    def make_reducer_backend(self, device):
        if device.type != "cuda":
            raise RuntimeError("the 'nccl'/'b200' backend reduces CUDA gradients only; this module lives on %s" % device)
        self._ensure_cuda(device.index)
        return B200ReducerBackend(self)

    # -- collectives ----------------------------------------------------------------------
    def _stream(self):
        return ctypes.c_void_p(torch.cuda.current_stream(self.device_index).cuda_stream)

    def _pull(self, root, nbytes, dst_ptr, blocks=8):
        st = self._stage()
        seg = st["seg"]
        n16 = (nbytes + 15) // 16
        src = _ptr_array([seg.ptr(r, st["pad"]) for r in range(self.size())])
        _check(_lib.cuda().pdl_pull_launch(self.rank(), self.size(), 0, root, n16, src, ctypes.c_void_p(dst_ptr), st["pads"],
                                           ctypes.c_void_p(st["state"].data_ptr()), blocks, self.timeout_s, self._stream()), "peer copy kernel")

    def broadcast(self, tensor, src):
        if not tensor.is_cuda:
            return self._ctl.broadcast(tensor, src)
        self._ensure_cuda(tensor.device.index)
        if self.size() == 1:
            return tensor
        st = self._stage()
        seg, half = st["seg"], st["half"]
        work = tensor.contiguous()
        flat = work.view(-1).view(torch.uint8)
        send = seg.tensor(st["pad"], half, torch.uint8)
        recv = seg.tensor(st["pad"] + half, half, torch.uint8)
        for off in range(0, flat.numel(), half):
            n = min(half, flat.numel() - off)
            if self.rank() == src:
                send[:n].copy_(flat[off:off + n])
            self._pull(src, n, recv.data_ptr(), blocks=min(32, max(1, n >> 16)))
            if self.rank() != src:
                flat[off:off + n].copy_(recv[:n])
        if work is not tensor:
            tensor.copy_(work)
        return tensor

    def allgather(self, tensor):
        if not tensor.is_cuda:
            return self._ctl.allgather(tensor)
        self._ensure_cuda(tensor.device.index)
        W = self.size()
        work = tensor.contiguous()
        if W == 1:
            return [work.clone()]
        st = self._stage()
        seg, half = st["seg"], st["half"]
        nbytes = work.numel() * work.element_size()
        send = seg.tensor(st["pad"], half, torch.uint8)
        recv = seg.tensor(st["pad"] + half, half, torch.uint8)
        flat = work.view(-1).view(torch.uint8)
        outs = [torch.empty(nbytes, dtype=torch.uint8, device=work.device) for _ in range(W)]
        chunk = half // W // 16 * 16  # bytes per rank per pass: W slots must fit the receive window
        for off in range(0, max(nbytes, 1), chunk):
            n = min(chunk, nbytes - off)
            slot = (n + 15) // 16 * 16
            send[:n].copy_(flat[off:off + n])
            self._pull(-1, slot, recv.data_ptr(), blocks=min(32, max(1, slot >> 14)))
            for r in range(W):
                outs[r][off:off + n].copy_(recv[r * slot:r * slot + n])
        return [o.view(work.dtype).view(work.shape) for o in outs]

    def allreduce(self, tensor, op=dist.ReduceOp.SUM, algorithm=None):
        if not tensor.is_cuda:
            return self._ctl.allreduce(tensor, op)
        self._ensure_cuda(tensor.device.index)
        W = self.size()
        if W == 1:
            return tensor
        if tensor.dtype != torch.float32 or op not in (dist.ReduceOp.SUM, dist.ReduceOp.AVG):
            # other dtypes / MIN / MAX / PRODUCT: gather every rank's tensor over peer memory (byte-level pull kernel) and reduce
            # the W copies locally in rank order -- stays on the device and on the stream, identical result on every rank
            stacked = torch.stack(self.allgather(tensor.detach()))
            if op == dist.ReduceOp.SUM:
                red = stacked.sum(0, dtype=tensor.dtype)
            elif op == dist.ReduceOp.AVG:
                red = stacked.sum(0, dtype=tensor.dtype) / W if tensor.is_floating_point() else torch.div(stacked.sum(0), W, rounding_mode="floor")
            elif op == dist.ReduceOp.MIN:
                red = stacked.amin(0)
            elif op == dist.ReduceOp.MAX:
                red = stacked.amax(0)
            elif op == dist.ReduceOp.PRODUCT:
                red = stacked.prod(0, dtype=tensor.dtype)
            else:
                raise ValueError("unsupported reduce op %r" % (op,))
            tensor.copy_(red.to(tensor.dtype))
            return tensor
        st = self._stage()
        seg, half, L = st["seg"], st["half"], _lib.cuda()
        work = tensor.contiguous()
        flat = work.view(-1)
        send = seg.tensor(st["pad"], half // 4)
        recv = seg.tensor(st["pad"] + half, half // 4)
        cap = half // 4 // 2  # floats per pass (second half of the send window is the ring staging area)
        scale = 1.0 if op == dist.ReduceOp.SUM else 1.0 / W
        ins = _ptr_array([seg.ptr(r, st["pad"]) for r in range(W)])
        outs = _ptr_array([seg.ptr(r, st["pad"] + half) for r in range(W)])
        rings = _ptr_array([seg.ptr(r, st["pad"] + cap * 4) for r in range(W)])
        push_row4 = self.PUSH_STAGE_BYTES // 16
        off = 0
        while off < flat.numel():
            algo = choose_algorithm(min(cap, flat.numel() - off) * 4, W, False, algorithm, table=False)
            n = min(push_row4 * 4 if algo == ALGOS["push"] else cap, flat.numel() - off)
            n4 = (n + 3) // 4
            send[:n].copy_(flat[off:off + n])
            if n4 * 4 > n:
                send[n:n4 * 4].zero_()
            blocks = L.pdl_allreduce_blocks(n4, W, algo)
            aux, row4 = rings, n4
            if algo == ALGOS["push"]:  # the group's own inbox: fixed rows, so launches of any length <= a row can share it
                if "push" not in st:
                    pseg = self.new_segment(int(L.pdl_allreduce_push_inbox_bytes(push_row4, W)), want_multicast=False)
                    st["push"] = _ptr_array([pseg.ptr(r) for r in range(W)])
                aux, row4 = st["push"], push_row4
            _check(L.pdl_allreduce_launch_ex(algo, self.rank(), W, 0, n4, row4, ins, outs, aux, st["pads"], None, None, None, None, None,
                                             ctypes.c_void_p(st["state"].data_ptr()), blocks, scale, self.timeout_s, self._stream()), "allreduce kernel")
            flat[off:off + n].copy_(recv[:n])
            off += n
        if work is not tensor:
            tensor.copy_(work)
        return tensor

    def barrier(self):
        if self.device_index is not None and self._staging is not None:
            torch.cuda.synchronize(self.device_index)
            self.check_device_errors()  # a host-side synchronisation point: the cheapest place to notice a dead peer
        self._ctl.barrier()

    def device_barrier(self):
        """Barrier executed by a kernel on the current stream (no host round trip)."""
        self._ensure_cuda()
        if self.size() > 1:
            st = self._stage()
            self._pull(0, 0, st["seg"].ptr(self.rank(), st["pad"] + st["half"]), blocks=1)

    def send(self, tensor, dst):
        host = tensor.detach().cpu() if tensor.is_cuda else tensor
        self._ctl.send(host, dst)

    def recv(self, tensor, src):
        if tensor.is_cuda:
            host = torch.empty(tensor.shape, dtype=tensor.dtype)
            self._ctl.recv(host, src)
            tensor.copy_(host)
            return tensor
        return self._ctl.recv(tensor, src)

    def check_device_errors(self, raise_on_error=True):
        """Failure detection for the device-side collectives: every peer wait inside a kernel is bounded
        (``PDL_KERNEL_TIMEOUT_S``, default 60 s) and a kernel that gives up sets the error word of its communicator state
        instead of spinning forever (a dead or hung rank must not wedge the GPU).  This reads those words (one small D2H
        copy per communicator, so it is called at host synchronisation points: ``barrier()``, checkpoints, group
        destruction, or explicitly) and raises.  The data produced by a timed-out kernel is garbage."""
        bad = 0
        word = _lib.cuda().pdl_comm_state_error_offset() // 4
        for ref in self._comm_states:
            st = ref()
            if st is not None and int(st.view(torch.int32)[word].item()) != 0:
                bad += 1
        if bad and raise_on_error:
            raise RuntimeError("[rank %d] %d device communicator(s) of group %r timed out waiting for a peer GPU (PDL_KERNEL_TIMEOUT_S=%gs): "
                               "a rank died or hung; gradients/parameters of this step are not valid" % (self.rank(), bad, self.name, self.timeout_s))
        return bad

    # control plane
    def _cpu_allgather(self, t):
        return self._ctl.allgather(t)

    def _cpu_broadcast(self, t, src):
        return self._ctl.broadcast(t, src)

    def destroy(self):
        failed = 0
        if self.device_index is not None:
            torch.cuda.synchronize(self.device_index)
            failed = self.check_device_errors(raise_on_error=False)
        try:
            self._ctl.barrier()
        except Exception:
            pass
        for seg in self._segments:
            seg.destroy()
        self._segments = []
        self._staging = None
        self._ctl.destroy()
        if failed:
            raise RuntimeError("[rank %d] group %r is being destroyed after %d of its device communicators lost a peer GPU; results of the "
                               "last steps are not valid" % (self.rank(), self.name, failed))


class B200ReducerBackend:
    """DDP's data plane on GPUs: buckets in symmetric memory, one fused kernel per bucket."""

    supports_fused_sgd = True

    def __init__(self, pg):
        self.pg = pg
        self.L = _lib.cuda()
        self.seg = None
        self.g_in = self.g_avg = None
        self.p_flat = self.m_flat = None
        self.total = 0
        self.collect_stats = os.environ.get("PDL_STATS", "0") == "1"
        self._exposed_ms = []
        self._launch_cache = {}
        self._hyper = {}
        self._last_done = None
        self._bw_end = None
        self._algo_used = {}
        self._attached = []
        self._reserved_channels = 0

    def allocate(self, total, dtype, device, bucket_ranges):
        if dtype != torch.float32:
            raise NotImplementedError("the NVLink reducer averages fp32 gradients (the reference trains in fp32); got %s" % dtype)
        self._launch_cache.clear()
        self._detach_fused_params()  # a re-layout moves every offset: params leave the flat arena first
        if self.seg is not None and total <= self.total:
            return
        pg, L = self.pg, self.L
        self.total = total
        self.pad_bytes = (L.pdl_signal_pad_bytes() + 4095) // 4096 * 4096
        self.area = (total * 4 + 4095) // 4096 * 4096
        self.seg = pg.new_segment(self.pad_bytes + 3 * self.area)
        self.state = pg.new_comm_state()
        self.g_in = self.seg.tensor(self.pad_bytes, total)
        self.g_avg = self.seg.tensor(self.pad_bytes + self.area, total)
        self.device = device
        W = pg.size()
        self._pads = _ptr_array([self.seg.ptr(r) for r in range(W)])

    # -- module buffers (BatchNorm running statistics) ------------------------------------------
    def make_buffer_sync(self, buffers):
        """DDP broadcasts rank 0's buffers before every forward (reference stack: two ncclBroadcast per forward for a
        ResNet).  Here the buffers are moved ONCE into a block of symmetric memory -- same offsets on every
        rank, every dtype, 16-byte aligned -- and the per-forward sync is a single kernel in which every replica pulls
        rank 0's block over NVLink into its own (flag round before: rank 0's values are final; flag round after: rank 0 may
        overwrite them).  No concatenation, no staging copies, no scatter.  Returns the callable, or None without buffers."""
        bufs = [b for b in buffers if b is not None and b.is_cuda]
        if not bufs:
            return None
        W, pg, L = self.pg.size(), self.pg, self.L
        offs, off = [], 0
        for b in bufs:
            offs.append(off)
            off = (off + b.numel() * b.element_size() + 15) // 16 * 16
        pad = (L.pdl_signal_pad_bytes() + 4095) // 4096 * 4096
        seg = pg.new_segment(pad + off, want_multicast=False)
        state = pg.new_comm_state()
        with torch.no_grad():
            for b, o in zip(bufs, offs):
                view = seg.tensor(pad + o, b.numel(), b.dtype).view(b.shape)
                view.copy_(b)
                b.data = view
        srcs = _ptr_array([seg.ptr(r, pad) for r in range(W)])
        pads = _ptr_array([seg.ptr(r) for r in range(W)])
        dst = ctypes.c_void_p(seg.ptr(pg.rank(), pad))
        n16 = off // 16
        blocks = min(32, max(1, off >> 15))
        keep = (seg, state, srcs, pads)

        def sync():
            stream = ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream)
            _check(L.pdl_pull_launch(pg.rank(), W, 0, 0, n16, keep[2], dst, keep[3], ctypes.c_void_p(keep[1].data_ptr()), blocks, pg.timeout_s, stream),
                   "buffer broadcast kernel")

        return sync

    # -- signal-pad channels ------------------------------------------------------------------
    # A channel is one epoch counter + one set of flag slots.  Two kernels may share a channel only if they can never be in
    # flight together, i.e. they are ordered on one stream.  DDP's own buckets all run on pg.comm_stream and share the
    # channels [0, shared); a caller that launches bucket kernels from any other stream takes a private channel from the top.
    def reserve_channel(self):
        top = self.L.pdl_comm_max_channels()
        if self._reserved_channels + 1 >= top:
            raise RuntimeError("no free signal-pad channel left (%d reserved of %d)" % (self._reserved_channels, top))
        self._reserved_channels += 1
        self._launch_cache.clear()
        return top - self._reserved_channels

    def _channel_of(self, bk):
        ch = getattr(bk, "channel", None)
        if ch is not None:
            return ch
        return bk.index % (self.L.pdl_comm_max_channels() - self._reserved_channels)

    # -- fused optimizer plumbing -----------------------------------------------------------
    def _detach_fused_params(self):
        with torch.no_grad():
            for opt, group, params in getattr(self, "_attached", []):
                for p in params:
                    p.data = p.data.clone()
                    st = opt.state.get(p)
                    if st is not None and st.get("momentum_buffer") is not None:
                        st["momentum_buffer"] = st["momentum_buffer"].clone()
                    opt.unbind_momentum_view(p)
        self._attached = []
        self._hyper = {}

    def attach_fused_params(self, opt, group, params, offsets):
        if any(p.dtype != torch.float32 or not p.is_cuda for p in params):
            return False
        if self.p_flat is None:
            self.p_flat = torch.zeros(self.total, dtype=torch.float32, device=self.device)
        mu = group["momentum"]
        if mu != 0 and self.m_flat is None:
            self.m_flat = torch.zeros(self.total, dtype=torch.float32, device=self.device)
        had_state = False
        with torch.no_grad():
            for p, off in zip(params, offsets):
                view = self.p_flat[off:off + p.numel()].view(p.shape)
                view.copy_(p.data)
                p.data = view
                if mu != 0:
                    mview = self.m_flat[off:off + p.numel()].view(p.shape)
                    old = opt.state[p].get("momentum_buffer") if p in opt.state else None
                    if old is not None:
                        mview.copy_(old)
                        had_state = True
                    opt.bind_momentum_view(p, mview)
        if not any(o is opt for o, _, _ in self._attached):
            opt.add_reload_listener(self._on_optimizer_reload)
        self._attached.append((opt, group, list(params)))
        return {"first": 0 if had_state else 1}

    def _on_optimizer_reload(self, first):
        """``optimizer.load_state_dict`` ran: upload every device-resident SgdHyper again (new learning rate etc.), with the
        first-step flag saying whether the restored state carries momentum history."""
        for entry in self._hyper.values():
            entry["vals"], entry["first"] = None, first

    def _hyper_for(self, bk):
        """Device-resident SgdHyper of a bucket, re-uploaded only when the optimizer's numbers change."""
        opt, gid = bk.fused_group
        group = opt.group_by_id(gid)
        key = bk.index
        entry = self._hyper.get(key)
        vals = (float(group["lr"]), float(group["momentum"]), float(group["dampening"]), float(group["weight_decay"]), int(bool(group["nesterov"])))
        if entry is None:
            dev = torch.zeros(8, dtype=torch.float32, device=self.device)
            entry = self._hyper[key] = {"dev": dev, "hosts": [], "vals": None, "first": getattr(bk, "first_flag", 1)}
        if entry["vals"] != vals:
            h = torch.zeros(8, dtype=torch.float32)
            h[0], h[1], h[2], h[3] = vals[0], vals[1], vals[2], vals[3]
            hi = h.view(torch.int32)
            hi[4] = vals[4]
            hi[5] = entry["first"]
            h = h.pin_memory()
            entry["hosts"] = entry["hosts"][-7:] + [h]  # keep in-flight pinned sources alive
            entry["dev"].copy_(h, non_blocking=True)
            entry["vals"] = vals
        entry["first"] = 0
        return entry["dev"]

    def bucket_launch_args(self, bk, world, algorithm):
        """(algo, blocks, ctypes pointer arrays) for a bucket; cached, also used by graph capture."""
        key = (bk.index, algorithm)
        c = self._launch_cache.get(key)
        if c is None:
            seg, W = self.seg, world
            off = bk.offset * 4
            n4 = (bk.numel + 3) // 4
            algo = choose_algorithm(bk.numel * 4, W, bool(seg.mc_ptr), algorithm)
            ring = [seg.ptr(r, self.pad_bytes + 2 * self.area + off) for r in range(W)]
            if algo == ALGOS["push"] and W > 1:
                # the bucket's own inbox ([2][W][n4] float4, collective allocation -- every rank resolves its buckets in the same
                # order): only with a private inbox does one flag round per launch suffice (csrc/cuda/allreduce.cu, PUSH)
                inbox = self.pg.new_segment(int(self.L.pdl_allreduce_push_inbox_bytes(n4, W)), want_multicast=False)
                ring = [inbox.ptr(r) for r in range(W)]
            c = {
                "algo": algo, "n4": n4, "blocks": choose_blocks(bk.numel * 4, W, algo, bool(seg.mc_ptr), algorithm),
                "in": _ptr_array([seg.ptr(r, self.pad_bytes + off) for r in range(W)]),
                "out": _ptr_array([seg.ptr(r, self.pad_bytes + self.area + off) for r in range(W)]),
                "ring": _ptr_array(ring),
                "mc_in": ctypes.c_void_p(seg.mc_ptr + self.pad_bytes + off) if seg.mc_ptr else None,
                "mc_out": ctypes.c_void_p(seg.mc_ptr + self.pad_bytes + self.area + off) if seg.mc_ptr else None,
                "channel": self._channel_of(bk),
            }
            self._launch_cache[key] = c
            self._algo_used[bk.index] = ALGO_NAMES[algo]
        return c

    def launch_bucket_kernel(self, bk, world, algorithm, stream_handle, fused):
        c = self.bucket_launch_args(bk, world, algorithm)
        param = mom = hyper = None
        if fused:
            off = bk.offset * 4
            param = ctypes.c_void_p(self.p_flat.data_ptr() + off)
            group = bk.fused_group[0].group_by_id(bk.fused_group[1])
            if group["momentum"] != 0:
                mom = ctypes.c_void_p(self.m_flat.data_ptr() + off)
            hyper = ctypes.c_void_p(self._hyper_for(bk).data_ptr())
        _check(self.L.pdl_allreduce_launch(c["algo"], self.pg.rank(), world, c["channel"], c["n4"], c["in"], c["out"], c["ring"], self._pads,
                                           c["mc_in"], c["mc_out"], param, mom, hyper, ctypes.c_void_p(self.state.data_ptr()), c["blocks"],
                                           1.0 / world, self.pg.timeout_s, ctypes.c_void_p(stream_handle)), "bucket allreduce kernel")

    def _bucket_handle(self, bk, world, algorithm):
        c = self.bucket_launch_args(bk, world, algorithm)
        h = c.get("handle")
        if h is None:
            h = ctypes.c_void_p()
            row4 = c["n4"]
            _check(self.L.pdl_bucket_create(c["algo"], self.pg.rank(), world, c["channel"], c["n4"], row4, c["in"], c["out"], c["ring"], self._pads,
                                            c["mc_in"], c["mc_out"], ctypes.c_void_p(self.state.data_ptr()), c["blocks"], 1.0 / world,
                                            self.pg.timeout_s, ctypes.byref(h)), "bucket handle")
            c["handle"] = h
        return h

    def reduce_bucket(self, bk, world, algorithm, fused=None):
        if fused is None:
            fused = bk.fused_group is not None
        if not self.collect_stats:
            # the hot path: ONE native call records "ready", orders the communication stream behind it, launches the fused
            # kernel there and records "done" (csrc/cuda/allreduce.cu: pdl_bucket_reduce)
            h = self._bucket_handle(bk, world, algorithm)
            param = mom = hyper = None
            if fused:
                off = bk.offset * 4
                param = ctypes.c_void_p(self.p_flat.data_ptr() + off)
                if bk.fused_group[0].group_by_id(bk.fused_group[1])["momentum"] != 0:
                    mom = ctypes.c_void_p(self.m_flat.data_ptr() + off)
                hyper = ctypes.c_void_p(self._hyper_for(bk).data_ptr())
            _check(self.L.pdl_bucket_reduce(h, param, mom, hyper, ctypes.c_void_p(torch.cuda.current_stream(self.device).cuda_stream),
                                            ctypes.c_void_p(self.pg.comm_stream.cuda_stream)), "bucket allreduce kernel")
            self._last_done = h
            if fused:
                bk.fused_group[0].mark_fused_done(bk.fused_params)
            return
        comm = self.pg.comm_stream
        cur = torch.cuda.current_stream(self.device)
        ready = torch.cuda.Event()
        ready.record(cur)
        comm.wait_event(ready)
        if fused is None:
            fused = bk.fused_group is not None
        with torch.cuda.stream(comm):
            self.launch_bucket_kernel(bk, world, algorithm, comm.cuda_stream, fused)
        done = torch.cuda.Event(enable_timing=self.collect_stats)
        done.record(comm)
        self._last_done = done
        if fused:
            opt = bk.fused_group[0]
            opt.mark_fused_done(bk.fused_params)

    def finish(self, buckets):
        self.pg.poll_errors()  # one host load: a lost peer reported by an earlier step's kernel raises here
        cur = torch.cuda.current_stream(self.device)
        if self.collect_stats:
            self._bw_end = torch.cuda.Event(enable_timing=True)
            self._bw_end.record(cur)
        if self._last_done is not None:
            if isinstance(self._last_done, ctypes.c_void_p):  # native handle: buckets run in order on one stream, the last "done" covers all
                _check(self.L.pdl_bucket_wait(self._last_done, ctypes.c_void_p(cur.cuda_stream)), "bucket wait")
            else:
                cur.wait_event(self._last_done)
                if self.collect_stats:
                    self._exposed_ms.append((self._bw_end, self._last_done))
        self._last_done = None

    def stats(self):
        out = {"backend": "b200", "algorithms": dict(self._algo_used), "multicast": bool(self.seg and self.seg.mc_ptr)}
        if self._exposed_ms:
            torch.cuda.synchronize(self.device)
            vals = [max(0.0, a.elapsed_time(b)) for a, b in self._exposed_ms[-50:]]
            out["exposed_comm_ms"] = sum(vals) / len(vals)
        return out
