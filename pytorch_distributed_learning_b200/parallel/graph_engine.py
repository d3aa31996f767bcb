"""Whole-iteration CUDA-graph capture for ANY model built from this package's (or torch's) layers.

The reference's loops are launch-bound: a ResNet-18 step at 32x32 is ~500 small kernels plus Python, DDP hooks
and the optimizer (example_launch.py:83-89).  Here the iteration -- forward, loss, backward with the DDP hooks
firing the fused allreduce+SGD kernels on the side stream, and whatever optimizer work is left -- is captured
once per input shape and replayed with a single launch.  Everything on the path is capture-safe by
construction: collectives keep their epochs in device memory, hyper-parameters live in device memory, scratch
buffers are grow-only and warmed up before capture.
"""
from __future__ import annotations

import torch


class GenericGraphEngine:
    WARMUP = 3

    def __init__(self, model, criterion, optimizer):
        self.model, self.criterion, self.optimizer = model, criterion, optimizer
        self.device = next(model.parameters()).device
        self.graphs = {}
        self._kernel_count = 0
        self._out = self._last = None
        self._grad_views = []
        self._ring = None  # pinned loss ring, created by the first step(pending=True)

    # -- helpers ---------------------------------------------------------------------------
    def _hyper_signature(self):
        keys = ("lr", "momentum", "dampening", "weight_decay", "nesterov")
        return tuple(tuple(g.get(k) for k in keys) for g in self.optimizer.param_groups)

    @staticmethod
    def supports(criterion):
        """Losses whose scalar result the engine can hand back: any module returning a 0-dim tensor works in principle; the
        transparent six-call mode only trusts this package's mean cross entropy."""
        from ..nn.modules import CrossEntropyLoss

        return isinstance(criterion, CrossEntropyLoss)

    def last_loss(self):
        """Device scalar holding the most recent step's loss (the graph's static buffer)."""
        return self._last["loss"]

    def grad_views(self):
        """(parameter, tensor that holds its averaged gradient after a replay) pairs."""
        return self._grad_views

    def updated_params(self):
        """Parameters whose update happens inside the captured iteration (all of them: the capture ends with optimizer.step())."""
        return [p for p, _ in self._grad_views]

    def _eager_step(self, x, y):
        from ..ops import tc_conv
        from . import lazy_step as _lazy

        with _lazy.suspend(), tc_conv.prepacked(self.model):  # one launch packs every conv weight for the tensor cores (fwd + dgrad copies)
            out = self.model(x)
            self._out = out.detach()
            loss = self.criterion(out, y)
            self.optimizer.zero_grad(set_to_none=True)
            loss.backward()
        self.optimizer.step()
        return loss.detach()

    def _refresh_fused_hyper(self):
        backend = getattr(self.model, "_backend", None)
        if backend is not None and hasattr(backend, "_hyper_for"):
            for bk in getattr(self.model, "_buckets", []):
                if bk.fused_group is not None:
                    backend._hyper_for(bk)

    def _build(self, images, labels):
        dev = self.device
        x = torch.empty(images.shape, dtype=images.dtype, device=dev)
        y = torch.empty(labels.shape, dtype=labels.dtype, device=dev)
        x.copy_(images)
        y.copy_(labels)
        # snapshot: the warm-up iterations must not count as training
        module = getattr(self.model, "module", self.model)
        params = [p for p in module.parameters()]
        saved_p = [p.detach().clone() for p in params]
        saved_b = [b.detach().clone() for b in module.buffers()]
        had_state = {id(p): (p in self.optimizer.state and "momentum_buffer" in self.optimizer.state[p]
                             and self.optimizer.state[p]["momentum_buffer"] is not None) for p in params}
        saved_m = {id(p): self.optimizer.state[p]["momentum_buffer"].detach().clone() for p in params if had_state[id(p)]}
        side = torch.cuda.Stream(device=dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            for _ in range(self.WARMUP):  # also triggers DDP's one-off bucket rebuild and the lazy optimizer fusion
                self._eager_step(x, y)
        torch.cuda.current_stream(dev).wait_stream(side)
        torch.cuda.synchronize(dev)
        with torch.no_grad():
            for p, s in zip(params, saved_p):
                p.copy_(s)
            for b, s in zip(module.buffers(), saved_b):
                b.copy_(s)
            for p in params:
                st = self.optimizer.state.get(p)
                if st is not None and st.get("momentum_buffer") is not None:
                    if had_state[id(p)]:
                        st["momentum_buffer"].copy_(saved_m[id(p)])
                    else:
                        st["momentum_buffer"].zero_()  # == "no history yet" for the default dampening of 0
        ddp = self.model if hasattr(self.model, "_rebuilt") else None
        if ddp is not None:
            ddp._rebuilt = True  # offsets are baked into the captured kernels
        self._refresh_fused_hyper()
        self.optimizer.zero_grad(set_to_none=True)
        torch.cuda.synchronize(dev)
        from .. import _lib

        L = _lib.cuda()
        before = L.pdl_launch_count()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            loss = self._eager_step(x, y)
        self._kernel_count = int(L.pdl_launch_count() - before)  # launcher calls of this library inside one iteration
        self._grad_views = [(p, p.grad) for p in params if p.grad is not None]
        return {"graph": g, "x": x, "y": y, "loss": loss, "out": self._out, "hyper": self._hyper_signature()}

    def step(self, images, labels, pending=False):
        """One iteration.  A host batch crosses PCIe on a copy stream into one of two staging buffers while the previous
        step computes; the graph's input is then refreshed by a device-to-device copy (38 MB for 64 ImageNet-sized images:
        ~25 us at HBM speed instead of ~0.7 ms of PCIe on the critical path).  ``pending=True`` returns a ``PendingLoss``."""
        key = (tuple(images.shape), images.dtype, tuple(labels.shape), labels.dtype)
        entry = self.graphs.get(key)
        if entry is not None and entry["hyper"] != self._hyper_signature() and not self._all_fused():
            entry = None  # un-fused optimizer kernels have the learning rate baked in: re-capture
        if entry is None:
            entry = self.graphs[key] = self._build(images, labels)
        self._refresh_fused_hyper()
        from ..ops import tc_conv
        from .pipeline import InputStager, refresh_inputs

        tc_conv.check_errors(self.device)  # a tensor-core kernel of an earlier replay gave up on its pipeline: raise (host load)
        pg = getattr(getattr(self.model, "_backend", None), "pg", None)
        if pg is not None and hasattr(pg, "poll_errors"):
            pg.poll_errors()               # a collective kernel of an earlier replay lost a peer

        if not images.is_cuda and not labels.is_cuda:
            stager = entry.get("stager")
            if stager is None:
                stager = entry["stager"] = InputStager(self.device)
            main = torch.cuda.current_stream(self.device)
            sx, sy, ready, slot = stager.stage(images, labels)
            main.wait_event(ready)
            refresh_inputs(entry["x"], entry["y"], sx, sy)
            stager.consumed(slot, main)
        else:
            refresh_inputs(entry["x"], entry["y"], images, labels)
        entry["graph"].replay()
        self._last = entry
        self.serial = getattr(self, "serial", 0) + 1
        if hasattr(self.model, "_iteration"):
            self.model._iteration += 1
        if pending:
            if self._ring is None:
                from .pipeline import LossRing

                self._ring = LossRing(self.device)
            return self._ring.publish(entry["loss"], torch.cuda.current_stream(self.device))[0]
        return entry["loss"]

    def outputs(self):
        """Model outputs of the most recent step (the graph's static buffer: the next step overwrites it)."""
        return self._last["out"]

    def _all_fused(self):
        buckets = getattr(self.model, "_buckets", None)
        return bool(buckets) and all(bk.fused_group is not None for bk in buckets)

    def kernels_per_step(self):
        return self._kernel_count
