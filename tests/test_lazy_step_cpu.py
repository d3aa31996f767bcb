"""The deferred objects behind the unchanged six-call loop (parallel/lazy_step.py), exercised on the CPU with a stand-in for
the wrapped model and the graph engine: what must happen is independent of CUDA -- defer while nothing needs a value, fall back
to the real forward pass the moment something does, never apply an update twice."""
import torch

import pytorch_distributed_learning_b200 as pdl
from pytorch_distributed_learning_b200.parallel import lazy_step
from pytorch_distributed_learning_b200.parallel.lazy_step import DeferredLoss, DeferredOutput


class _Pending:
    def __init__(self, v):
        self.v = v

    def item(self):
        return self.v


class _Engine:
    """Plays the whole-iteration graph: records that it ran, exposes logits / loss / grads like the real engines."""

    def __init__(self, model, opt):
        self.model, self.optimizer, self.serial, self.steps = model, opt, 0, 0
        self._logits = None

    def step(self, images, labels, pending=False):
        with torch.no_grad():
            self._logits = self.model(images)
            self._loss = torch.nn.functional.cross_entropy(self._logits, labels)
        self.serial += 1
        self.steps += 1
        return _Pending(float(self._loss))

    def outputs(self):
        return self._logits

    def last_loss(self):
        return self._loss

    def grad_views(self):
        return [(p, torch.zeros_like(p)) for p in self.model.parameters()]

    def updated_params(self):
        return list(self.model.parameters())


class _FakeDDP:
    device = torch.device("cpu")

    def __init__(self, model, engine):
        self.module, self.engine, self.forwards = model, engine, 0
        self._graph_step, self._graph_engines = "auto", {}

    def _forward_now(self, images):
        self.forwards += 1
        with lazy_step.suspend():
            return self.module(images)

    def _graph_engine(self, criterion):
        return self.engine


def _setup():
    torch.manual_seed(0)
    model = torch.nn.Linear(4, 3)
    opt = pdl.optim.SGD(model.parameters(), lr=0.1)
    ddp = _FakeDDP(model, None)
    ddp.engine = _Engine(model, opt)
    return model, opt, ddp, pdl.nn.CrossEntropyLoss(), torch.randn(5, 4), torch.randint(0, 3, (5,))


def test_deferred_loop_replays_the_engine_and_skips_the_second_update():
    model, opt, ddp, crit, x, y = _setup()
    out = DeferredOutput(ddp, x)
    loss = crit(out, y)
    assert isinstance(loss, DeferredLoss) and ddp.forwards == 0
    opt.zero_grad()
    before = [p.detach().clone() for p in model.parameters()]
    loss.backward()
    assert ddp.engine.steps == 1 and ddp.forwards == 0
    for p in model.parameters():
        assert p.grad is not None  # the averaged-gradient views are visible again after zero_grad()
    opt.step()  # everything was marked as updated by the "graph": no second update
    assert all(torch.equal(a, b.detach()) for a, b in zip(before, model.parameters()))
    assert abs(loss.item() - float(torch.nn.functional.cross_entropy(model(x), y))) < 1e-6
    assert out.argmax(dim=1).shape == (5,) and torch.max(out, 1)[1].shape == (5,)  # the engine's logits, cloned
    assert ddp.forwards == 0
    assert "{:.4f}".format(loss.item())


def test_using_the_outputs_first_falls_back_to_autograd():
    model, opt, ddp, crit, x, y = _setup()
    out = DeferredOutput(ddp, x)
    assert out.shape == (5, 3) and ddp.forwards == 1  # needs the real tensor: the forward pass runs now, with autograd
    loss = crit(out, y)
    opt.zero_grad()
    loss.backward()
    assert ddp.engine.steps == 0, "an iteration whose outputs were already computed must not replay the graph"
    assert model.weight.grad is not None and model.weight.grad.abs().sum() > 0
    w0 = model.weight.detach().clone()
    opt.step()
    assert not torch.equal(w0, model.weight.detach())  # the ordinary optimizer step did the update


def test_other_criteria_and_backward_arguments_use_real_tensors():
    model, opt, ddp, crit, x, y = _setup()
    out = DeferredOutput(ddp, x)
    loss = torch.nn.functional.cross_entropy(out, y)  # torch's functional form: __torch_function__ materialises
    assert isinstance(loss, torch.Tensor) and ddp.forwards == 1
    out2 = DeferredOutput(ddp, x)
    l2 = crit(out2, y)
    l2.backward(retain_graph=True)  # arguments the graph cannot honour: autograd path
    assert ddp.engine.steps == 0 and ddp.forwards == 2
    out3 = DeferredOutput(ddp, x)
    weighted = pdl.nn.CrossEntropyLoss(label_smoothing=0.1)(out3, y)  # not the loss the graphed step computes
    assert isinstance(weighted, torch.Tensor)


def test_stale_outputs_are_refused_not_silently_wrong():
    model, opt, ddp, crit, x, y = _setup()
    out = DeferredOutput(ddp, x)
    loss = crit(out, y)
    loss.backward()
    out_b = DeferredOutput(ddp, x)
    crit(out_b, y).backward()  # a later step reuses the graph's logits buffer
    try:
        out.argmax(dim=1)
        raised = False
    except RuntimeError as e:
        raised = "later step" in str(e)
    assert raised
    assert loss.item() == loss.item()  # the loss of the older step is still in the pinned ring
