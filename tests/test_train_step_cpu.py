"""TrainStep on a CPU model: the eager fallback; step_async returns a value .item() works on (no GPU plumbing involved)."""
import copy

import torch

import pytorch_distributed_learning_b200 as pdl


def test_cpu_train_step_and_step_async_agree():
    torch.manual_seed(0)
    a = pdl.models.ConvNet()
    b = copy.deepcopy(a)
    crit = pdl.nn.CrossEntropyLoss()
    sa = pdl.TrainStep(a, crit, pdl.optim.SGD(a.parameters(), lr=0.05))
    sb = pdl.TrainStep(b, crit, pdl.optim.SGD(b.parameters(), lr=0.05))
    assert not sa.graphed and sa.kernels_per_step() is None
    for i in range(3):
        x, y = torch.randn(4, 1, 28, 28), torch.randint(0, 10, (4,))
        la, lb = sa(x, y), sb.step_async(x, y)
        assert float(la.item()) == float(lb.item())
    for p, q in zip(a.parameters(), b.parameters()):
        assert torch.equal(p, q)


def test_pending_loss_bookkeeping_without_a_gpu():
    """PendingLoss reads its own ring slot after its own event and refuses to read a slot that has been reused."""
    from pytorch_distributed_learning_b200.parallel.pipeline import PendingLoss

    class _Event:
        def __init__(self):
            self.waited = 0

        def synchronize(self):
            self.waited += 1

        def query(self):
            return True

    class _Ring:
        DEPTH = 4

        def __init__(self):
            self.host, self.events, self.serial = torch.arange(4, dtype=torch.float32) * 1.5, [_Event() for _ in range(4)], 0

    ring = _Ring()
    handles = []
    for _ in range(4):
        handles.append(PendingLoss(ring, ring.serial % ring.DEPTH, ring.serial))
        ring.serial += 1
    assert [h.item() for h in handles] == [0.0, 1.5, 3.0, 4.5] and float(handles[2]) == 3.0
    assert ring.events[2].waited == 2 and handles[0].ready()
    ring.serial += 1  # a fifth step reuses slot 0
    try:
        handles[0].item()
    except RuntimeError as e:
        assert "expired" in str(e)
    else:
        raise AssertionError("expired handle was read")
    assert handles[1].item() == 1.5
