import pytest
import torch
import torch.distributed as tdist

from pytorch_distributed_learning_b200.models import ConvNet, resnet18, resnet50
from pytorch_distributed_learning_b200.parallel.ddp import compute_bucket_assignment_by_size

MiB = 1024 * 1024


def _sizes(model):
    return [p.numel() * p.element_size() for p in model.parameters()]


@pytest.mark.parametrize("factory", [ConvNet, lambda: resnet18(num_classes=10), resnet50])
@pytest.mark.parametrize("limits", [[MiB, 25 * MiB], [1 << 62], [4096, 8192, MiB]])
def test_matches_torch_assignment(factory, limits):
    model = factory()
    params = list(reversed(list(model.parameters())))
    ours = compute_bucket_assignment_by_size([p.numel() * 4 for p in params], limits)
    ref, _ = tdist._compute_bucket_assignment_by_size(params, limits)
    assert ours == [list(b) for b in ref]


def test_reference_bucket_bytes():
    # steady-state message sizes the reference's stack puts on the wire
    conv = list(reversed(_sizes(ConvNet())))
    b = compute_bucket_assignment_by_size(conv, [MiB, 25 * MiB])
    assert [sum(conv[i] for i in bk) for bk in b] == [454696]
    r50 = list(reversed(_sizes(resnet50())))
    b = compute_bucket_assignment_by_size(r50, [MiB, 25 * MiB])
    assert sum(sum(r50[i] for i in bk) for bk in b) == 102228128
    assert len(b) == 5


def test_kinds_never_share_a_bucket():
    b = compute_bucket_assignment_by_size([10, 10, 10, 10], [1000], kinds=[0, 1, 0, 1])
    assert b == [[0, 2], [1, 3]]


def test_param_counts():
    assert sum(p.numel() for p in ConvNet().parameters()) == 113674
    assert sum(p.numel() for p in resnet18(num_classes=10).parameters()) == 11181642
    assert sum(p.numel() for p in resnet50().parameters()) == 25557032
