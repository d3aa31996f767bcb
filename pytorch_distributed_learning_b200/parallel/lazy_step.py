"""The reference's six-call loop, unchanged, on the graphed step.

    outputs = model(images);  loss = criterion(outputs, labels)          # launch_dist.py:86-87
    optimizer.zero_grad();  loss.backward();  optimizer.step()            # launch_dist.py:90-92

Run call by call, a 0.12 ms iteration costs ~1 ms of Python, autograd and launch overhead (measured: about the same as the
reference's own stack).  Nothing in those five lines needs a result before ``loss.backward()``, so on a B200 the wrapped
model does not run its forward pass when it is called: ``DistributedDataParallel.forward`` hands back a ``DeferredOutput``,
this package's ``CrossEntropyLoss`` turns it into a ``DeferredLoss``, and ``loss.backward()`` replays the WHOLE iteration
-- forward, loss, backward, gradient exchange, SGD -- as the one CUDA graph ``TrainStep`` uses.  ``optimizer.step()`` then
finds its parameters already updated, exactly as on the fused eager path.

Everything else degrades to ordinary eager execution instead of failing: touching the outputs before ``backward()``
(``outputs.argmax(1)``, ``torch.max(outputs, 1)``, printing, indexing ...) runs the real forward pass then and there, a
criterion other than this package's ``CrossEntropyLoss`` receives real logits, ``backward(retain_graph=...)`` or a
custom ``gradient`` takes the autograd path.  After ``backward()`` the outputs and the loss read the graph's results
(``outputs.argmax(1)`` and ``loss.item()`` in the reference's ResNet loop, example_launch.py:91-93).

Opt out with ``DistributedDataParallel(..., graph_step=False)`` or ``PDL_TRANSPARENT_GRAPH=0``.
"""
from __future__ import annotations

import contextlib
import os
import threading

import torch

_state = threading.local()


def suspended():
    return getattr(_state, "depth", 0) > 0


@contextlib.contextmanager
def suspend():
    """Inside this block wrapped models run their forward pass immediately (used by the engines' own eager/capture passes)."""
    _state.depth = getattr(_state, "depth", 0) + 1
    try:
        yield
    finally:
        _state.depth -= 1


def enabled_by_default():
    return os.environ.get("PDL_TRANSPARENT_GRAPH", "1") != "0"


def _real(x):
    if isinstance(x, _Deferred):
        return x._tensor()
    if isinstance(x, (list, tuple)):
        return type(x)(_real(v) for v in x)
    if isinstance(x, dict):
        return {k: _real(v) for k, v in x.items()}
    return x


class _Deferred:
    """Stands in for a tensor that has not been computed yet.  Any use -- a torch function, an attribute, an operator --
    produces the real tensor first (``_tensor()``) and then behaves like it."""

    @classmethod
    def __torch_function__(cls, func, types, args=(), kwargs=None):
        return func(*_real(args), **_real(kwargs or {}))

    def __getattr__(self, name):
        if name.startswith("__") and name.endswith("__"):
            raise AttributeError(name)
        return getattr(self._tensor(), name)

    def __repr__(self):
        return repr(self._tensor())

    def __len__(self):
        return len(self._tensor())

    def __iter__(self):
        return iter(self._tensor())

    def __getitem__(self, idx):
        return self._tensor()[_real(idx)]

    def __bool__(self):
        return bool(self._tensor())

    def __float__(self):
        return float(self._tensor())

    def __int__(self):
        return int(self._tensor())

    def __format__(self, spec):
        return format(self._tensor(), spec)


def _binary(name):
    def op(self, other):
        return getattr(self._tensor(), name)(_real(other))

    op.__name__ = name
    return op


for _n in ("add", "radd", "sub", "rsub", "mul", "rmul", "truediv", "rtruediv", "pow", "matmul", "eq", "ne", "lt", "le", "gt", "ge", "neg"):
    if _n == "neg":
        setattr(_Deferred, "__neg__", lambda self: -self._tensor())
    else:
        setattr(_Deferred, "__%s__" % _n, _binary("__%s__" % _n))
_Deferred.__hash__ = object.__hash__


class DeferredOutput(_Deferred):
    def __init__(self, ddp, images):
        self._ddp, self._images = ddp, images
        self._value = None      # the real output, once something asked for it
        self._loss = None       # the DeferredLoss built on top of this output, if any
        self._eager = False     # True once the real forward pass ran with autograd (the graph is then not used)

    def _tensor(self):
        if self._value is None:
            loss = self._loss
            if loss is not None and loss._engine is not None:  # the graph ran: its logits buffer holds this step's outputs
                if loss._engine.serial != loss._serial:
                    raise RuntimeError("the outputs of this training step were first used after a later step had run; the graphed "
                                       "step keeps only the most recent step's outputs (use them before the next loss.backward(), "
                                       "or construct DistributedDataParallel with graph_step=False)")
                self._value = loss._engine.outputs().detach().clone()
            else:
                self._value = self._ddp._forward_now(self._images)
                self._eager = True
        return self._value


class DeferredLoss(_Deferred):
    def __init__(self, output, labels, criterion):
        self._out, self._labels, self._criterion = output, labels, criterion
        self._engine = None     # set when backward() replayed the graph
        self._serial = -1       # the engine's step counter right after that replay
        self._pending = None    # PendingLoss of that replay (pinned-host ring)
        self._value = None
        output._loss = self

    # -- the one call that does the work ---------------------------------------------------------------------------
    def backward(self, *args, **kwargs):
        out = self._out
        engine = None
        if not args and not kwargs and not out._eager and self._engine is None:
            engine = out._ddp._graph_engine(self._criterion)
        if engine is None:
            return self._tensor().backward(*args, **kwargs)
        ddp = out._ddp
        try:
            self._pending = engine.step(out._images, self._labels, pending=True)
        except Exception:
            ddp._graph_step = False  # e.g. an op that cannot be captured: stay on the eager path from now on
            ddp._graph_engines.clear()
            raise
        self._engine, self._serial = engine, engine.serial
        opt = engine.optimizer
        for p, g in engine.grad_views():
            if p.grad is None:
                p.grad = g  # optimizer.zero_grad() came in between; the averaged gradients are where they always are
        opt.mark_fused_done(engine.updated_params())  # optimizer.step() must not apply the update a second time
        return None

    def item(self):
        if self._pending is not None:
            return self._pending.item()
        return self._tensor().item()

    def __float__(self):
        return float(self.item())

    def detach(self):
        return self

    def _tensor(self):
        if self._value is None:
            if self._pending is not None:
                if self._engine.serial == self._serial:  # still the most recent step: copy the device scalar, no host round trip
                    self._value = self._engine.last_loss().detach().clone()
                else:  # a later step reused the buffer; the pinned-host ring still has this step's value
                    self._value = torch.tensor(self._pending.item(), dtype=torch.float32, device=self._out._ddp.device)
            else:
                with suspend():
                    self._value = self._criterion(self._out._tensor(), self._labels)
        return self._value
