"""DistributedDataParallel: replica-per-process data parallelism with bucketed gradient averaging.

Call shape kept from the reference: ``DistributedDataParallel(model, device_ids=[g],
output_device=g)`` (launch_dist.py:61, example_mp.py:53, example_launch.py:29) and
``DistributedDataParallel(model, device_ids=[g])`` (mpspawn_dist.py:68).

What happens, B200-first rather than a mirror of torch's reducer:
  * constructor: parameter shapes are checked across ranks, then rank 0's parameters and
    buffers are broadcast so every replica starts identical (the reference relies on this:
    its ResNet scripts never seed);
  * gradients live in ONE flat arena laid out bucket by bucket in reverse registration order
    (≈ the order backward produces them).  On GPUs the arena is *symmetric memory*: the same
    offsets on every rank, mapped into every peer, so the allreduce kernel reads the peers'
    buckets straight over NVLink -- there is no copy-out, ``param.grad`` becomes a view of the
    averaged arena;
  * a bucket is launched on a side stream the moment its last gradient arrives (autograd
    post-accumulate hooks), so communication overlaps the rest of backward;
  * if the parameters belong to this package's ``optim.SGD`` the averaging kernel also applies
    the SGD update (fused 1/W scale + weight decay + momentum + nesterov), and
    ``optimizer.step()`` only waits for the side stream;
  * the first backward records the real gradient arrival order; if it differs from the assumed
    one the buckets are rebuilt once, with rank 0's order, before the second forward.
"""
from __future__ import annotations

import ctypes
import os
import weakref
from contextlib import contextmanager

import torch
import torch.nn as nn
from torch.autograd import Variable

from .. import _lib
from ..utils import tracing as _tracing
from . import distributed as dist
from .reducer import ReducerCore

_MiB = 1024 * 1024
_ALIGN_ELEMS = 4  # every parameter starts on a 16-byte boundary inside the arena (float4 access)


def compute_bucket_assignment_by_size(sizes_bytes, limits, kinds=None):
    """C++ bucket assignment (csrc/host/bucket.cpp). Returns list[list[int]] of tensor indices."""
    n = len(sizes_bytes)
    if n == 0:
        return []
    sizes = (ctypes.c_int64 * n)(*[int(s) for s in sizes_bytes])
    kind_arr = (ctypes.c_int32 * n)(*([int(k) for k in kinds] if kinds is not None else [0] * n))
    lim = (ctypes.c_int64 * len(limits))(*[int(x) for x in limits])
    out = (ctypes.c_int32 * n)()
    nb = _lib.host().pdl_bucket_assign(sizes, kind_arr, n, lim, len(limits), out)
    buckets = [[] for _ in range(nb)]
    for i in range(n):
        buckets[out[i]].append(i)
    return buckets


class _Bucket:
    __slots__ = ("index", "param_ids", "offset", "numel", "launched", "fused_group", "fused_params", "first_flag", "grads", "channel")

    def __init__(self, index, param_ids, offset, numel):
        self.index, self.param_ids, self.offset, self.numel = index, param_ids, offset, numel
        # signal-pad channel (epoch counter + flag slots).  None = assigned by the backend: DDP launches every bucket on the ONE
        # communication stream, where sharing a channel is safe because the kernels are stream-ordered.  Anything launched from
        # another stream must own its channel (B200ReducerBackend.reserve_channel).
        self.channel = None
        self.launched = False  # pending counts and the launch order live in the C++ reducer core (reducer.py)
        self.fused_group, self.fused_params, self.first_flag = None, (), 1
        self.grads = {}


class GradBucket:
    """What a communication hook receives (torch.distributed.GradBucket's accessors)."""

    def __init__(self, index, flat, is_last, grads, params):
        self._index, self._flat, self._is_last, self._grads, self._params = index, flat, is_last, grads, params

    def index(self):
        return self._index

    def buffer(self):
        return self._flat

    def is_last(self):
        return self._is_last

    def gradients(self):
        return list(self._grads)

    def parameters(self):
        return list(self._params)

    def set_buffer(self, tensor):
        self._flat.copy_(tensor.reshape(-1))


def _align(n, a=_ALIGN_ELEMS):
    return (n + a - 1) // a * a


class DistributedDataParallel(nn.Module):
    def __init__(self, module, device_ids=None, output_device=None, dim=0, broadcast_buffers=True, process_group=None,
                 bucket_cap_mb=25, find_unused_parameters=False, check_reduction=False, gradient_as_bucket_view=True,
                 static_graph=False, algorithm=None, first_bucket_mb=1, fused_optimizer=None, graph_step=None):
        """``fused_optimizer`` (default on; ``PDL_FUSED_OPTIMIZER=0`` or ``False`` turns it off): when the parameters belong to
        this package's ``optim.SGD`` the bucket kernel applies the update itself, i.e. the weights change DURING
        ``loss.backward()`` and ``optimizer.step()`` has nothing left to do.  Anything placed between ``backward()`` and
        ``step()`` -- gradient clipping, edits of ``.grad``, skipping the step on a NaN -- needs ``fused_optimizer=False``:
        the kernel then only averages, ``.grad`` holds the averaged gradient and ``step()`` performs the whole update, exactly
        like the reference's loop (launch_dist.py:91-92)."""
        super().__init__()
        # graph_step: run the reference's unchanged six-call loop on the whole-iteration CUDA graph (parallel/lazy_step.py).
        # None = automatic for the model families this package ships (ConvNet, ResNet) on a GPU; True = also for any other
        # model built from capturable ops; False (or PDL_TRANSPARENT_GRAPH=0) = every call runs when it is made.
        from . import lazy_step as _lazy

        self._graph_step = ("auto" if _lazy.enabled_by_default() else False) if graph_step is None else bool(graph_step)
        self._graph_engines = {}
        if fused_optimizer is None:
            fused_optimizer = os.environ.get("PDL_FUSED_OPTIMIZER", "1") != "0"
        self.fused_optimizer = bool(fused_optimizer)
        self.process_group = process_group if process_group is not None else dist._default()
        params = [(n, p) for n, p in module.named_parameters() if p.requires_grad]
        if not params:
            raise RuntimeError("DistributedDataParallel is not needed when a module doesn't have any parameter that requires a gradient.")
        devices = {p.device for _, p in params}
        if len(devices) != 1:
            raise ValueError("DistributedDataParallel requires all parameters of the module on one device, got %s" % sorted(map(str, devices)))
        self.device = next(iter(devices))
        if self.device.type == "cuda":
            if device_ids is not None:
                if len(device_ids) != 1:
                    raise ValueError("device_ids can only be None or contain a single element.")
                dev = device_ids[0]
                dev_index = dev.index if isinstance(dev, torch.device) else int(dev)
                if dev_index != self.device.index:
                    raise ValueError("device_ids=[%d] but the module's parameters live on %s" % (dev_index, self.device))
            self.device_ids = [self.device.index]
            if output_device is None:
                output_device = self.device.index
            self.output_device = output_device.index if isinstance(output_device, torch.device) else int(output_device)
        else:
            if device_ids or output_device is not None:
                raise ValueError("device_ids and output_device must be None for a CPU module")
            self.device_ids, self.output_device = None, None
        self.module = module
        self.dim = dim
        self.broadcast_buffers = broadcast_buffers
        self.find_unused_parameters = find_unused_parameters
        self.bucket_bytes_cap = int(bucket_cap_mb * _MiB)
        self.first_bucket_bytes = int(first_bucket_mb * _MiB)
        self.algorithm = algorithm or os.environ.get("PDL_ALGO") or None
        self.require_backward_grad_sync = True
        self._param_names = [n for n, _ in params]
        self._params = [p for _, p in params]
        self._world = self.process_group.size()
        self._iteration = 0
        self._core = None  # C++ reducer state machine, created with the first bucket layout
        self._rebuilt = False
        self._callback_queued = False
        self._comm_hook, self._hook_results = None, []
        self._stats = {"buckets_launched": 0, "fused_buckets": 0, "rebuilds": 0}

        self._verify_params_across_processes()
        self._sync_module_states()
        self._backend = self.process_group.make_reducer_backend(self.device) if hasattr(self.process_group, "make_reducer_backend") else _CpuReducerBackend(self.process_group)
        self._build_buckets(list(reversed(range(len(self._params)))))
        self._buffer_sync = None
        if self.broadcast_buffers and self._world > 1 and hasattr(self._backend, "make_buffer_sync"):
            self._buffer_sync = self._backend.make_buffer_sync(list(self.module.buffers()))
        self._hook_handles = [p.register_post_accumulate_grad_hook(self._make_hook(i)) for i, p in enumerate(self._params)]

    # ------------------------------------------------------------------ construction
    def _verify_params_across_processes(self):
        meta = [(n, tuple(p.shape), str(p.dtype)) for n, p in zip(self._param_names, self._params)]
        everyone = self.process_group.allgather_object(meta)
        for r, other in enumerate(everyone):
            if other != everyone[0]:
                raise RuntimeError("DDP expects the same model on every rank: rank %d's parameter list differs from rank 0's "
                                   "(%d vs %d tensors, first mismatch: %s)" %
                                   (r, len(other), len(everyone[0]), next((a for a, b in zip(other, everyone[0]) if a != b), "length")))

    def _sync_module_states(self, chunk_bytes=250 * _MiB):
        tensors = [p.data for p in self.module.parameters()] + [b for b in self.module.buffers()]
        self._broadcast_coalesced(tensors, chunk_bytes)

    def _broadcast_coalesced(self, tensors, chunk_bytes=250 * _MiB):
        by_dtype = {}
        for t in tensors:
            by_dtype.setdefault(t.dtype, []).append(t)
        for dtype, group in by_dtype.items():
            chunk, size = [], 0
            for t in group + [None]:
                if t is None or (chunk and size + t.numel() * t.element_size() > chunk_bytes):
                    if chunk:
                        flat = torch.cat([c.reshape(-1) for c in chunk])
                        self.process_group.broadcast(flat, 0)
                        off = 0
                        for c in chunk:
                            c.copy_(flat[off:off + c.numel()].view_as(c))
                            off += c.numel()
                    chunk, size = [], 0
                if t is not None:
                    chunk.append(t)
                    size += t.numel() * t.element_size()

    def _build_buckets(self, order):
        """order: parameter indices in expected gradient-arrival order."""
        ps = self._params
        dtypes = sorted({p.dtype for p in ps}, key=str)
        if len(dtypes) != 1:
            raise NotImplementedError("DistributedDataParallel: mixed parameter dtypes %s are not supported; cast the model to one dtype" % dtypes)
        self._dtype = dtypes[0]
        sizes = [ps[i].numel() * ps[i].element_size() for i in order]
        assignment = compute_bucket_assignment_by_size(sizes, [self.first_bucket_bytes, self.bucket_bytes_cap])
        self._buckets, self._bucket_of, self._offset_of = [], {}, {}
        off = 0
        for b, members in enumerate(assignment):
            start = off
            ids = [order[j] for j in members]
            for i in ids:
                self._bucket_of[i] = b
                self._offset_of[i] = off
                off = _align(off + ps[i].numel())
            self._buckets.append(_Bucket(b, ids, start, off - start))
        self._total = off
        self._backend.allocate(self._total, self._dtype, self.device, [(bk.offset, bk.numel) for bk in self._buckets])
        g_in, g_avg = self._backend.g_in, self._backend.g_avg
        self._in_views = [g_in[self._offset_of[i]:self._offset_of[i] + p.numel()].view_as(p) for i, p in enumerate(ps)]
        self._avg_views = [g_avg[self._offset_of[i]:self._offset_of[i] + p.numel()].view_as(p) for i, p in enumerate(ps)]
        self._order = list(order)
        layout = [self._bucket_of[i] for i in range(len(ps))]
        if self._core is None:
            self._core = ReducerCore(len(ps), layout)
        else:
            self._core.set_layout(len(ps), layout)
        self._fused_checked = False

    def bucket_sizes_bytes(self):
        """Unpadded bytes per bucket (what the reference's stack would put on the wire)."""
        return [sum(self._params[i].numel() * self._params[i].element_size() for i in bk.param_ids) for bk in self._buckets]

    # ------------------------------------------------------------------ forward
    def forward(self, *inputs, **kwargs):
        if self._graph_step is not False and self._can_defer(inputs, kwargs):
            from .lazy_step import DeferredOutput

            return DeferredOutput(self, inputs[0])
        return self._forward_traced(*inputs, **kwargs)

    def _can_defer(self, inputs, kwargs):
        from . import lazy_step as _lazy

        if (kwargs or len(inputs) != 1 or not isinstance(inputs[0], torch.Tensor) or self.device.type != "cuda" or not self.training
                or not torch.is_grad_enabled() or not self.require_backward_grad_sync or self._comm_hook is not None
                or not self.fused_optimizer or not getattr(self._backend, "supports_fused_sgd", False) or _lazy.suspended()):
            return False
        if self._graph_step is True:
            return True
        from ..models.convnet import ConvNet
        from ..models.resnet import ResNet

        return isinstance(self.module, (ConvNet, ResNet))

    def _forward_now(self, images):
        """The real forward pass of a deferred call (something needed the outputs before ``loss.backward()``)."""
        from . import lazy_step as _lazy

        with _lazy.suspend():
            return self._forward_traced(images)

    def _graph_engine(self, criterion):
        """The whole-iteration graph for (this model, ``criterion``, the optimizer that owns the parameters), or None when the
        combination is not one the graphed step covers -- the caller then runs eagerly, and so does every later call."""
        if self._graph_step is False:
            return None
        key = id(criterion)
        if key in self._graph_engines:
            return self._graph_engines[key][0]
        from ..optim.sgd import lookup_optimizer

        owners = {id(o[0]): o[0] for o in (lookup_optimizer(p) for p in self._params) if o is not None}
        engine = None
        if len(owners) == 1 and all(lookup_optimizer(p) is not None for p in self._params):
            opt = next(iter(owners.values()))
            from ..ops.convnet_engine import ConvNetEngine

            try:
                engine = ConvNetEngine.try_create(self, criterion, opt)
                if engine is None:
                    from .graph_engine import GenericGraphEngine

                    if GenericGraphEngine.supports(criterion):
                        engine = GenericGraphEngine(self, criterion, opt)
            except (RuntimeError, ValueError, NotImplementedError) as e:
                # e.g. an optimizer group the fused update does not cover: the loop simply runs call by call
                from ..utils.logging import debug

                debug("transparent graph step disabled: %s" % e)
                engine = None
        self._graph_engines[key] = (engine, criterion)  # keeps the criterion alive: ids are only unique among live objects
        if engine is None:
            self._graph_step = False
        return engine

    def _forward_traced(self, *inputs, **kwargs):
        if _tracing.enabled():
            with _tracing.trace_range("pdl.DistributedDataParallel.forward", iteration=self._iteration):
                return self._forward(*inputs, **kwargs)
        return self._forward(*inputs, **kwargs)

    def _forward(self, *inputs, **kwargs):
        if torch.is_grad_enabled() and self.require_backward_grad_sync:
            self._maybe_rebuild_buckets()
            self._prepare_for_backward()
        else:
            self._remove_grad_sinks()  # a local (no_sync / no-grad) pass must not write into the exchange arena
        if self.broadcast_buffers and self.require_backward_grad_sync:
            if self._buffer_sync is not None:
                self._buffer_sync()  # ONE kernel: every replica pulls rank 0's buffer block over NVLink (K14)
            else:
                bufs = list(self.module.buffers())
                if bufs and self._world > 1:
                    self._broadcast_coalesced(bufs)
        if self.device.type == "cuda":
            inputs = tuple(self._to_device(x) for x in inputs)
            kwargs = {k: self._to_device(v) for k, v in kwargs.items()}
        return self.module(*inputs, **kwargs)

    def _to_device(self, x):
        if isinstance(x, torch.Tensor) and x.device != self.device:
            return x.to(self.device, non_blocking=True)
        return x

    def _prepare_for_backward(self):
        for bk in self._buckets:
            bk.launched = False
            bk.grads.clear()
        self._core.prepare(record_arrival=self._iteration == 0)  # pending counts, ready flags, launch cursor (C++)
        self._callback_queued = False
        if self.device.type == "cuda" and self._comm_hook is None:
            if not self._fused_checked:
                self._resolve_fused_optimizers()  # may move the parameters into the flat arena: do it before forward saves them
            self._install_grad_sinks()

    def _install_grad_sinks(self):
        """Tell the native backward functions where each parameter's gradient belongs (its view of the symmetric arena): their
        kernels write it there and autograd adopts the alias as ``param.grad`` -- no gradient tensor of its own, no copy
        into the bucket (``ops.functional.grad_buffer``).  Removed again when backward ends."""
        from ..ops import functional as NF

        self._sink_keys = [p.data_ptr() for p in self._params]
        NF.install_grad_sinks({k: v for k, v in zip(self._sink_keys, self._in_views)})

    def _remove_grad_sinks(self):
        keys = getattr(self, "_sink_keys", None)
        if keys:
            from ..ops import functional as NF

            NF.remove_grad_sinks(keys)
            self._sink_keys = None

    def _maybe_rebuild_buckets(self):
        if self._rebuilt or self._iteration != 1:
            return
        self._rebuilt = True
        seen = list(dict.fromkeys(self._core.arrival()))
        seen += [i for i in self._order if i not in set(seen)]
        order = self.process_group.broadcast_object(seen, 0)
        if order != self._order:
            self._stats["rebuilds"] += 1
            old_avg = [v.clone() for v in self._avg_views]
            self._build_buckets(order)
            for i, p in enumerate(self._params):  # keep any grads the user still holds valid
                self._avg_views[i].copy_(old_avg[i])
                if p.grad is not None:
                    p.grad = self._avg_views[i]

    # ------------------------------------------------------------------ backward
    def _make_hook(self, i):
        ref = weakref.ref(self)

        def hook(param):
            self_ = ref()
            if self_ is not None:
                self_._on_grad_ready(i, param)

        return hook

    def _on_grad_ready(self, i, param):
        if not self.require_backward_grad_sync:
            return
        if not self._callback_queued:
            self._callback_queued = True
            Variable._execution_engine.queue_callback(self._finalize_backward)
        try:
            launchable = self._core.mark_ready(i)  # buckets complete now AND next in the fixed launch order
        except RuntimeError:
            raise RuntimeError("parameter %r received a second gradient in one backward pass; DDP reduces each parameter once per "
                               "iteration (use no_sync() for gradient accumulation)" % self._param_names[i]) from None
        self._buckets[self._bucket_of[i]].grads[i] = param.grad
        for b in launchable:
            self._launch_bucket(self._buckets[b])

    def _launch_bucket(self, bk):
        if _tracing.enabled():
            with _tracing.trace_range("pdl.bucket", index=self._buckets.index(bk), params=len(bk.param_ids), fused=bk.fused_group is not None):
                return self._launch_bucket_impl(bk)
        return self._launch_bucket_impl(bk)

    def _launch_bucket_impl(self, bk):
        srcs, dsts = [], []
        unused = []
        for i in bk.param_ids:
            g = bk.grads.get(i)
            view = self._in_views[i]
            if g is None:  # parameter took no part in this backward: contributes zeros
                view.zero_()
                unused.append(i)
            elif g.data_ptr() != view.data_ptr():
                srcs.append(g)
                dsts.append(view)
        if dsts:
            torch._foreach_copy_(dsts, srcs)
        if self._comm_hook is not None:
            self._run_comm_hook(bk)
            return
        if not self._fused_checked:
            self._resolve_fused_optimizers()
        # A parameter without a gradient must not be touched by the optimizer (torch skips ``grad is None``: no weight decay,
        # no momentum step).  The fused kernel updates whole buckets, so such a bucket is only averaged this iteration and
        # optimizer.step() updates the parameters that did receive a gradient.
        fused = bk.fused_group is not None and not unused
        self._backend.reduce_bucket(bk, self._world, self.algorithm, fused=fused) if bk.fused_group is not None else \
            self._backend.reduce_bucket(bk, self._world, self.algorithm)
        skip = set(unused)
        for i in bk.param_ids:
            if i not in skip:
                self._params[i].grad = self._avg_views[i]
        bk.launched = True
        bk.grads.clear()
        self._stats["buckets_launched"] += 1
        if fused:
            self._stats["fused_buckets"] += 1

    def _finalize_backward(self):
        self._callback_queued = False
        self._remove_grad_sinks()
        if not self.require_backward_grad_sync:
            return
        for b in self._core.finalize():  # some parameters produced no gradient this iteration: they contribute zeros
            self._launch_bucket(self._buckets[b])
        if self._comm_hook is not None:
            self._finish_comm_hook()
        else:
            self._backend.finish(self._buckets)
        self._iteration += 1

    # ------------------------------------------------------------------ communication hook
    def register_comm_hook(self, state, hook):
        """``hook(state, bucket) -> tensor | future`` takes over the reduction of every gradient bucket (same contract
        as torch's DDP: the bucket holds the UN-divided local gradients, the hook returns -- directly, or through an
        object with ``wait()`` / ``value()`` -- the tensor that becomes the gradients, e.g. the allreduced sum divided
        by the world size).  With a hook installed the built-in bucket kernel and its fused SGD update are bypassed;
        the optimizer's ``step()`` then performs the whole update.  May be registered once, before the first backward."""
        if not callable(hook):
            raise TypeError("communication hook must be callable")
        if self._comm_hook is not None:
            raise RuntimeError("register_comm_hook can only be called once")
        if self._iteration > 0:
            raise RuntimeError("register_comm_hook must be called before the first backward pass")
        self._comm_hook = (state, hook)
        self._hook_results = []

    def _run_comm_hook(self, bk):
        state, hook = self._comm_hook
        flat = self._backend.g_in[bk.offset:bk.offset + bk.numel]
        out = hook(state, GradBucket(bk.index, flat, bk.index == len(self._buckets) - 1,
                                     [self._in_views[i] for i in bk.param_ids], [self._params[i] for i in bk.param_ids]))
        self._hook_results.append((bk, out))
        bk.launched = True
        bk.grads.clear()
        self._stats["buckets_launched"] += 1

    def _finish_comm_hook(self):
        for bk, out in self._hook_results:
            if not torch.is_tensor(out):
                res = out.wait() if hasattr(out, "wait") else out
                if res is None or not torch.is_tensor(res):
                    res = out.value() if hasattr(out, "value") else res
                if isinstance(res, (list, tuple)):
                    res = res[0]
                out = res
            if not torch.is_tensor(out) or out.numel() != bk.numel:
                raise RuntimeError("communication hook must return a tensor with the bucket's %d elements" % bk.numel)
            dst = self._backend.g_avg[bk.offset:bk.offset + bk.numel]
            if out.data_ptr() != dst.data_ptr():
                dst.copy_(out.reshape(-1))
            for i in bk.param_ids:
                self._params[i].grad = self._avg_views[i]
        self._hook_results = []

    # ------------------------------------------------------------------ fused optimizer
    def _resolve_fused_optimizers(self):
        """Look up (lazily: the optimizer may be created before or after the wrap, reference
        launch_dist.py:59-61 vs example_launch.py:29,64) which buckets can take the fused update."""
        self._fused_checked = True
        if not self._backend.supports_fused_sgd or not self.fused_optimizer:
            return
        from ..optim.sgd import lookup_optimizer

        for bk in self._buckets:
            owners = {lookup_optimizer(self._params[i]) for i in bk.param_ids}
            if len(owners) != 1:
                continue
            owner = next(iter(owners))
            if owner is None:
                continue
            opt, group = owner
            members = [self._params[i] for i in bk.param_ids]
            info = opt.attach_bucket(self, bk, members, [self._offset_of[i] for i in bk.param_ids], group)
            if info:
                bk.fused_group, bk.fused_params, bk.first_flag = (opt, group), members, info.get("first", 1)

    # ------------------------------------------------------------------ API parity
    @contextmanager
    def no_sync(self):
        """Skip gradient averaging inside the block (local accumulation into ``.grad``)."""
        old = self.require_backward_grad_sync
        self.require_backward_grad_sync = False
        try:
            yield
        finally:
            self.require_backward_grad_sync = old

    def check_health(self):
        """Raise if a bucket kernel of this group gave up waiting for a peer GPU (dead / hung rank); see
        ``B200ProcessGroup.check_device_errors``.  Costs a device synchronisation: call it where you already synchronise
        (when you print the loss, before a checkpoint).  No-op on the CPU backend, whose collectives raise by themselves."""
        check = getattr(self.process_group, "check_device_errors", None)
        if check is not None and self.device.type == "cuda":
            torch.cuda.synchronize(self.device)
            check()

    def get_step_stats(self):
        stats = dict(self._stats)
        stats["bucket_bytes"] = self.bucket_sizes_bytes()
        stats["iterations"] = self._iteration
        stats.update(self._backend.stats())
        return stats

    def state_dict(self, *args, **kwargs):
        return super().state_dict(*args, **kwargs)

    def train(self, mode=True):
        super().train(mode)
        return self


class _CpuReducerBackend:
    """Reducer data plane for the CPU ("gloo") group: in-place allreduce, then scale by 1/W."""

    supports_fused_sgd = False

    def __init__(self, pg):
        self.pg = pg
        self.g_in = self.g_avg = None

    def allocate(self, total, dtype, device, bucket_ranges):
        self.g_in = torch.zeros(total, dtype=dtype, device=device)
        self.g_avg = self.g_in

    def reduce_bucket(self, bk, world, algorithm):
        flat = self.g_in[bk.offset:bk.offset + bk.numel]
        if world > 1:
            self.pg.allreduce(flat, dist.ReduceOp.SUM)
            flat.mul_(1.0 / world)

    def finish(self, buckets):
        pass

    def stats(self):
        return {"backend": "cpu"}
