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
