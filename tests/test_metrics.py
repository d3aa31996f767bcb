"""utils.stats.RunningMetrics: running loss / top-1 accuracy (reference example_launch.py:91-96 semantics)."""
import pytest
import torch

import pytorch_distributed_learning_b200 as pdl


def _reference(outs, tgts, losses):
    train_loss = correct = total = 0
    for o, t, l in zip(outs, tgts, losses):  # the reference's per-step bookkeeping
        train_loss += l.item()
        total += t.size(0)
        correct += torch.eq(o.argmax(dim=1), t).sum().item()
    return train_loss / len(outs), correct / total, total


def test_running_metrics_cpu_matches_reference_bookkeeping():
    g = torch.Generator().manual_seed(0)
    outs = [torch.randn(37, 10, generator=g) for _ in range(5)]
    tgts = [torch.randint(0, 10, (37,), generator=g) for _ in range(5)]
    losses = [torch.rand((), generator=g) for _ in range(5)]
    m = pdl.utils.stats.RunningMetrics()
    for o, t, l in zip(outs, tgts, losses):
        m.update(o, t, l)
    got, want = m.read(), _reference(outs, tgts, losses)
    assert got[0] == pytest.approx(want[0], rel=1e-6) and got[1] == pytest.approx(want[1]) and got[2] == want[2]
    m.reset()
    assert m.read() == (0.0, 0.0, 0)


@pytest.mark.gpu
@pytest.mark.parametrize("N,C", [(100, 10), (64, 1000), (5, 3), (257, 33)])
def test_running_metrics_kernel_matches_torch(N, C):
    dev = torch.device("cuda", 0)
    g = torch.Generator().manual_seed(N * 1000 + C)
    outs = [torch.randn(N, C, generator=g) for _ in range(4)]
    outs[1][:, : C // 2 + 1] = outs[1][:, :1]  # ties: the first maximum must win
    outs[2][0, C - 1] = float("nan")  # a NaN counts as the maximum
    tgts = [torch.randint(0, C, (N,), generator=g) for _ in range(4)]
    tgts[1][: N // 2] = 0
    tgts[2][0] = C - 1
    losses = [torch.rand((), generator=g) for _ in range(4)]
    m = pdl.utils.stats.RunningMetrics(dev)
    for o, t, l in zip(outs, tgts, losses):
        m.update(o.to(dev), t.to(dev), l.to(dev))
    m.update(outs[0].to(dev), tgts[0].to(dev))  # without a loss: only the counts move
    got = m.read()
    want = _reference(outs + [outs[0]], tgts + [tgts[0]], losses + [torch.zeros(())])
    assert got[2] == want[2] and got[1] == pytest.approx(want[1], abs=1e-12)
    assert got[0] == pytest.approx(want[0], rel=1e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("which", ["convnet", "generic"])
def test_train_step_outputs_feed_the_metrics(which):
    dev = torch.device("cuda", 0)
    torch.manual_seed(5)
    if which == "convnet":
        model, shape = pdl.models.ConvNet().to(dev), (1, 28, 28)
    else:
        model, shape = pdl.models.resnet18(num_classes=10).to(dev), (3, 32, 32)
    step = pdl.TrainStep(model, pdl.nn.CrossEntropyLoss(), pdl.optim.SGD(model.parameters(), lr=0.01))
    m = pdl.utils.stats.RunningMetrics(dev)
    x, y = torch.randn((16,) + shape, device=dev), torch.randint(0, 10, (16,), device=dev)
    ref_loss = ref_hits = 0.0
    for _ in range(3):
        loss = step(x, y)
        out = step.outputs()
        assert out.shape == (16, 10)
        m.update(out, y, loss)
        ref_loss += loss.item()
        ref_hits += (out.argmax(1) == y).sum().item()
    got = m.read()
    assert got[2] == 48 and got[1] == pytest.approx(ref_hits / 48) and got[0] == pytest.approx(ref_loss / 3, rel=1e-6)
    # the logits are those of the loss that was reported
    assert torch.nn.functional.cross_entropy(out, y).item() == pytest.approx(loss.item(), rel=1e-4)
