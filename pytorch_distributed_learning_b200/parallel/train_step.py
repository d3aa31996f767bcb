"""``TrainStep``: one training iteration as a single call.

``step = TrainStep(ddp_model, criterion, optimizer); loss = step(images, labels)`` is
equivalent to the reference's loop body (launch_dist.py:82-92):

    images, labels -> device;  out = model(images);  loss = criterion(out, labels)
    optimizer.zero_grad();  loss.backward();  optimizer.step()

On a B200, when the wrapped module is this package's ``models.ConvNet`` trained with
``nn.CrossEntropyLoss`` and ``optim.SGD``, the whole iteration -- H2D copy of the batch, fused
conv+bias+ReLU+pool forward kernels, loss, backward kernels writing gradients straight into
the symmetric bucket, the peer-to-peer allreduce fused with 1/W averaging and the SGD update --
is captured ONCE into a CUDA graph and replayed with one launch per step (the reference spends
its time in ~40 launches + Python + NCCL latency per step).  Any other model / device runs the
same sequence eagerly.
"""
from __future__ import annotations

import torch

from ..utils import tracing as _tracing


class TrainStep:
    def __init__(self, model, criterion, optimizer, use_graph=True):
        self.model, self.criterion, self.optimizer = model, criterion, optimizer
        self.engine = None
        self._ring = self._out = None
        if use_graph:
            self.engine = self._try_native_engine()

    def _try_native_engine(self):
        module = getattr(self.model, "module", self.model)
        try:
            p = next(module.parameters())
        except StopIteration:
            return None
        if not p.is_cuda:
            return None
        from ..ops.convnet_engine import ConvNetEngine

        engine = ConvNetEngine.try_create(self.model, self.criterion, self.optimizer)
        if engine is None:  # any other model: capture the eager iteration (hooks, side stream and all) as one graph
            from .graph_engine import GenericGraphEngine

            engine = GenericGraphEngine(self.model, self.criterion, self.optimizer)
        return engine

    @property
    def graphed(self):
        return self.engine is not None

    def __call__(self, images, labels):
        if _tracing.enabled():
            with _tracing.trace_range("pdl.TrainStep", graphed=self.graphed, batch=int(images.shape[0])):
                return self._step(images, labels)
        return self._step(images, labels)

    def _step(self, images, labels):
        if self.engine is not None:
            return self.engine.step(images, labels)
        return self._eager(images, labels)

    def step_async(self, images, labels):
        """Like ``step(images, labels)`` but returns a ``PendingLoss``: the scalar is copied to pinned host memory right
        behind the step, and ``.item()`` / ``float()`` waits for that copy only.  Read it one step late --

            pending = None
            for images, labels in loader:            # pinned host batches
                cur = step.step_async(images, labels)
                if pending is not None: log(pending.item())
                pending = cur

        -- and neither the next batch's H2D copy nor the loss read-back ever leaves the GPU idle (the reference's loop
        reads ``loss.item()`` every 100th step for the same reason, launch_dist.py:93-99).  A handle stays valid for
        ``LossRing.DEPTH`` further steps."""
        if self.engine is not None:
            return self.engine.step(images, labels, pending=True)
        loss = self._eager(images, labels)
        if not loss.is_cuda:
            return loss  # CPU training: the value is already on the host; .item() works the same way
        if self._ring is None:
            from .pipeline import LossRing

            self._ring = LossRing(loss.device)
        return self._ring.publish(loss, torch.cuda.current_stream(loss.device))[0]

    def _eager(self, images, labels):
        from . import lazy_step as _lazy

        dev = next(self.model.parameters()).device
        if images.device != dev:
            images = images.to(dev, non_blocking=True)
            labels = labels.to(dev, non_blocking=True)
        with _lazy.suspend():  # use_graph=False means call by call: no transparent graph either
            out = self.model(images)
            self._out = out.detach()
            loss = self.criterion(out, labels)
            self.optimizer.zero_grad()
            loss.backward()
            self.optimizer.step()
        return loss.detach()

    def outputs(self):
        """Model outputs (logits) of the most recent step, for metrics such as ``utils.stats.RunningMetrics``.  Under a
        graph this is a static buffer that the next step overwrites; consume it (on the same stream) before stepping again."""
        if self.engine is not None:
            return self.engine.outputs()
        return self._out

    def kernels_per_step(self):
        return self.engine.kernels_per_step() if self.engine is not None else None
