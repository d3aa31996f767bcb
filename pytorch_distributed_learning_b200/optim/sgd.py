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
    for group in opt.param_groups:
        for p in group["params"]:
            if p is param:
                return opt, id(group)
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

    def add_param_group(self, param_group):
        super().add_param_group(param_group)
        for p in self.param_groups[-1]["params"]:
            _registry[id(p)] = self

    def group_by_id(self, gid):
        for g in self.param_groups:
            if id(g) == gid:
                return g
        raise KeyError("parameter group no longer exists")

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
