"""SGD whose update can be fused into the gradient-averaging kernel.

Same constructor and update rule as the optimizer the reference uses --
``SGD(params, 1e-4)`` (launch_dist.py:59) and ``SGD(params, lr=0.02, momentum=0.9,
weight_decay=1e-4, nesterov=True)`` (example_launch.py:64-70):

    g += wd * p;  buf = g (first step) | mu * buf + (1 - damp) * g;  g = g + mu * buf (nesterov) | buf;  p -= lr * g

It subclasses ``torch.optim.SGD`` so ``state_dict()`` / ``load_state_dict()`` / ``param_groups``
/ LR schedulers keep their layout.  When the parameters are wrapped by this package's
DistributedDataParallel on GPUs, the bucket allreduce kernel applies the update itself and
``step()`` only skips what was already done; anything not covered (no DDP, ``no_sync``
iterations, CPU) goes through the stand-alone multi-tensor kernel (GPU) or torch's CPU path.
"""
from __future__ import annotations

import weakref

import torch

_registry = weakref.WeakValueDictionary()  # id(param) -> optimizer


def lookup_optimizer(param):
    """-> (optimizer, param_group) if ``param`` is owned by one of our SGD instances."""
    opt = _registry.get(id(param))
    if opt is None:
        return None
    for gi, group in enumerate(opt.param_groups):
        for p in group["params"]:
            if p is param:
                return opt, gi  # the group's INDEX: load_state_dict() replaces the dict objects, positions stay
    return None


class SGD(torch.optim.SGD):
    def __init__(self, params, lr=1e-3, momentum=0.0, dampening=0.0, weight_decay=0.0, nesterov=False, *, maximize=False, **kwargs):
        kwargs.pop("foreach", None)
        kwargs.pop("fused", None)
        super().__init__(params, lr=lr, momentum=momentum, dampening=dampening, weight_decay=weight_decay, nesterov=nesterov,
                         maximize=maximize, **kwargs)
        for group in self.param_groups:
            for p in group["params"]:
                _registry[id(p)] = self
        self._fused_done = set()  # ids of params already updated by a fused bucket this iteration
        self._attached = []  # (weakref(ddp), bucket, group_id)
        self._native_tables = {}
        self._momentum_views = {}  # id(param) -> view of a flat momentum arena that fused kernels update in place
        self._reload_listeners = []  # weak callables(first_step) run after load_state_dict (device-resident hyper-parameters)

    def __setstate__(self, state):
        """Unpickling / ``copy.deepcopy`` restores torch's three fields only (``Optimizer.__getstate__``): the copy starts with
        empty fusion bookkeeping -- it owns different parameter objects, so nothing of the original's arenas applies to it."""
        super().__setstate__(state)
        self.__dict__.setdefault("_fused_done", set())
        self.__dict__.setdefault("_attached", [])
        self.__dict__.setdefault("_native_tables", {})
        self.__dict__.setdefault("_momentum_views", {})
        self.__dict__.setdefault("_reload_listeners", [])
        for group in self.param_groups:
            for p in group["params"]:
                _registry[id(p)] = self

    def add_param_group(self, param_group):
        super().add_param_group(param_group)
        for p in self.param_groups[-1]["params"]:
            _registry[id(p)] = self

    def group_by_id(self, gid):
        """Parameter group by position (``gid`` as returned by ``lookup_optimizer``)."""
        try:
            return self.param_groups[gid]
        except (IndexError, TypeError):
            raise KeyError("parameter group %r no longer exists" % (gid,)) from None

    # -- flat-arena bookkeeping (called by the fused back ends) ---------------------------
    def bind_momentum_view(self, param, view):
        """``view`` (a slice of a flat arena that the fused kernels read and write) IS this parameter's momentum buffer."""
        self._momentum_views[id(param)] = view
        self.state[param]["momentum_buffer"] = view

    def unbind_momentum_view(self, param):
        self._momentum_views.pop(id(param), None)

    def add_reload_listener(self, bound_method):
        self._reload_listeners.append(weakref.WeakMethod(bound_method))

    def load_state_dict(self, state_dict):
        """torch's ``load_state_dict`` replaces every group dict and every state tensor.  The fused kernels keep updating the
        flat arenas, so the loaded momentum is copied INTO the arena views (and the views put back as the state), and whoever
        mirrors hyper-parameters in device memory is told to upload them again -- a resumed run continues exactly like an
        uninterrupted one (reference convention: rank-0 ``state_dict`` checkpoints)."""
        super().load_state_dict(state_dict)
        have_history = False
        with torch.no_grad():
            for group in self.param_groups:
                for p in group["params"]:
                    view = self._momentum_views.get(id(p))
                    st = self.state.get(p)
                    loaded = st.get("momentum_buffer") if st is not None else None
                    if loaded is not None:
                        have_history = True
                    if view is None:
                        continue
                    if loaded is None:
                        view.zero_()
                    elif loaded is not view:
                        view.copy_(loaded)
                    self.state[p]["momentum_buffer"] = view
        self._fused_done = set()
        alive = []
        for ref in self._reload_listeners:
            fn = ref()
            if fn is not None:
                fn(0 if have_history else 1)
                alive.append(ref)
        self._reload_listeners = alive

    # -- called by DistributedDataParallel ---------------------------------------------
    def attach_bucket(self, ddp, bucket, params, offsets, group_id):
        group = self.group_by_id(group_id)
        if group.get("maximize", False):
            return False
        info = ddp._backend.attach_fused_params(self, group, params, offsets)
        if not info:
            return False
        self._attached.append((weakref.ref(ddp), bucket, group_id))
        return info

    def mark_fused_done(self, params):
        self._fused_done.update(id(p) for p in params)

    # -- torch.optim API ---------------------------------------------------------------
    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        done = self._fused_done
        for group in self.param_groups:
            todo = [p for p in group["params"] if p.grad is not None and id(p) not in done]
            if not todo:
                continue
            if todo[0].is_cuda and todo[0].dtype == torch.float32 and not group.get("maximize", False):
                self._native_step(group, todo)
            else:
                self._torch_step(group, todo)
        self._fused_done = set()
        return loss

    def _momentum_buffers(self, group, params):
        if group["momentum"] == 0:
            return [None] * len(params)  # like torch: no state entry is created without momentum
        return [self.state[p].get("momentum_buffer") for p in params]

    def _torch_step(self, group, params):
        from torch.optim.sgd import sgd as functional_sgd

        grads = [p.grad for p in params]
        bufs = self._momentum_buffers(group, params)
        functional_sgd(params, grads, bufs, weight_decay=group["weight_decay"], momentum=group["momentum"], lr=group["lr"],
                       dampening=group["dampening"], nesterov=group["nesterov"], maximize=group.get("maximize", False),
                       has_sparse_grad=False, foreach=False)
        if group["momentum"] != 0:
            for p, b in zip(params, bufs):
                self.state[p]["momentum_buffer"] = b

    def _native_step(self, group, params):
        from ..ops import sgd as sgd_ops

        mu = group["momentum"]
        first = []
        for p in params:
            if mu != 0 and self.state[p].get("momentum_buffer") is None:
                self.state[p]["momentum_buffer"] = torch.zeros_like(p, memory_format=torch.contiguous_format)
                first.append(True)
            else:
                first.append(False)
        bufs = [self.state[p].get("momentum_buffer") for p in params] if mu != 0 else [None] * len(params)
        sgd_ops.multi_tensor_sgd(params, [p.grad for p in params], bufs, first, lr=group["lr"], momentum=mu,
                                 dampening=group["dampening"], weight_decay=group["weight_decay"], nesterov=group["nesterov"])
