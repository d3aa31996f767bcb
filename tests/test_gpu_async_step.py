"""TrainStep.step_async: host batches staged on a copy stream through two alternating input slots and the loss read
back through a pinned ring must train exactly like the blocking path (same kernels, same order -> same numbers)."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu

import pytorch_distributed_learning_b200 as pdl  # noqa: E402


def _batches(n, shape, classes, batch):
    g = torch.Generator().manual_seed(7)
    xs = [torch.randn((batch,) + shape, generator=g).pin_memory() for _ in range(n)]
    ys = [torch.randint(0, classes, (batch,), generator=g).pin_memory() for _ in range(n)]
    return xs, ys


def _twin_steps(make_model, opt_kw):
    dev = torch.device("cuda", 0)
    torch.manual_seed(11)
    a = make_model().to(dev)
    b = copy.deepcopy(a)
    crit = pdl.nn.CrossEntropyLoss()
    sa = pdl.TrainStep(a, crit, pdl.optim.SGD(a.parameters(), **opt_kw))
    sb = pdl.TrainStep(b, crit, pdl.optim.SGD(b.parameters(), **opt_kw))
    return a, b, sa, sb


@pytest.mark.parametrize("which", ["convnet", "generic"])
def test_step_async_matches_blocking_steps(which):
    dev = torch.device("cuda", 0)
    if which == "convnet":
        make, shape, batch, kw, tol = pdl.models.ConvNet, (1, 28, 28), 24, dict(lr=0.05), 1e-6
    else:
        make, shape, batch, kw, tol = (lambda: pdl.models.resnet18(num_classes=10)), (3, 32, 32), 8, dict(lr=0.05, momentum=0.9), 1e-4
    a, b, sa, sb = _twin_steps(make, kw)
    xs, ys = _batches(7, shape, 10, batch)
    want = [float(sb(x.to(dev), y.to(dev)).item()) for x, y in zip(xs, ys)]  # blocking, device-resident inputs
    got, prev = [], None
    for x, y in zip(xs, ys):  # read one step late, as a training loop would
        cur = sa.step_async(x, y)
        assert isinstance(cur, pdl.PendingLoss)
        if prev is not None:
            got.append(prev.item())
        prev = cur
    got.append(float(prev))
    assert prev.ready()
    assert got == pytest.approx(want, rel=tol, abs=tol), (got, want)
    for (n, p), (_, q) in zip(a.named_parameters(), b.named_parameters()):
        assert torch.allclose(p, q, rtol=tol, atol=tol), n
    # the two entry points can be mixed: a blocking step on a host batch, then a device batch, land on the same numbers
    la = float(sa(xs[0], ys[0]).item())
    lb = float(sb(xs[0].to(dev), ys[0].to(dev)).item())
    assert la == pytest.approx(lb, rel=tol)
    la = float(sa(xs[1].to(dev), ys[1].to(dev)).item())
    lb = float(sb(xs[1], ys[1]).item())
    assert la == pytest.approx(lb, rel=tol)


def test_pending_loss_expires_instead_of_returning_a_later_step():
    dev = torch.device("cuda", 0)
    model = pdl.models.ConvNet().to(dev)
    step = pdl.TrainStep(model, pdl.nn.CrossEntropyLoss(), pdl.optim.SGD(model.parameters(), lr=0.01))
    xs, ys = _batches(2, (1, 28, 28), 10, 16)
    first = step.step_async(xs[0], ys[0])
    depth = first._ring.DEPTH
    kept = [step.step_async(xs[i % 2], ys[i % 2]) for i in range(depth)]
    with pytest.raises(RuntimeError, match="expired"):
        first.item()
    assert all(torch.isfinite(torch.tensor(h.item())) for h in kept[1:])


@pytest.mark.parametrize("n0,n1,off", [(313600, 800, 0), (1000, 7, 0), (4099, 33, 4), (16, 16, 1), (0, 24, 0)])
def test_copy_pair_kernel(n0, n1, off):
    import ctypes

    from pytorch_distributed_learning_b200 import _lib

    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(n0 + n1)
    s0 = torch.randint(0, 255, (n0 + off,), dtype=torch.uint8, generator=g).to(dev)
    s1 = torch.randint(0, 255, (n1 + off,), dtype=torch.uint8, generator=g).to(dev)
    d0 = torch.zeros(n0 + off + 16, dtype=torch.uint8, device=dev)
    d1 = torch.zeros(n1 + off + 16, dtype=torch.uint8, device=dev)
    P = ctypes.c_void_p
    rc = _lib.cuda().pdl_copy_pair(P(d0.data_ptr() + off), P(s0.data_ptr() + off), n0, P(d1.data_ptr() + off), P(s1.data_ptr() + off), n1,
                                   P(torch.cuda.current_stream().cuda_stream))
    assert rc == 0
    assert torch.equal(d0[off:off + n0], s0[off:off + n0]) and torch.equal(d1[off:off + n1], s1[off:off + n1])
    assert d0[:off].sum() == 0 and d0[off + n0:].sum() == 0 and d1[:off].sum() == 0 and d1[off + n1:].sum() == 0  # nothing outside the ranges


def test_cuda_prefetcher_delivers_every_batch_in_order():
    dev = torch.device("cuda", 0)
    ds = pdl.utils.data.SyntheticCIFAR10(1000)
    loader = torch.utils.data.DataLoader(ds, batch_size=96, shuffle=False, pin_memory=True)
    pre = pdl.utils.data.CudaPrefetcher(loader, dev)
    seen = 0
    acc = torch.zeros((), device=dev, dtype=torch.float64)
    for (x, y), (rx, ry) in zip(pre, loader):
        assert x.is_cuda and y.is_cuda and x.shape == rx.shape
        # keep the GPU busy between batches so the next copy really overlaps queued work
        acc += (x.double() @ torch.ones(x.shape[-1], device=dev, dtype=torch.float64)).sum()
        assert torch.equal(x.cpu(), rx) and torch.equal(y.cpu(), ry)
        seen += x.shape[0]
    assert seen == 1000 and len(pre) == 11 and torch.isfinite(acc)
    assert sum(b[0].shape[0] for b in pre) == 1000  # a second epoch works
