"""Skip connection routed through the opening convolution's autograd node (conv(x, skip=True)): the skip gradient is
added in the dgrad kernel's epilogue.  Must give the same gradients as the plain formulation."""
import copy

import pytest
import torch

pytestmark = pytest.mark.gpu

import pytorch_distributed_learning_b200 as pdl  # noqa: E402


@pytest.fixture(autouse=True)
def _tf32_convs():
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True
    yield
    torch.backends.cudnn.allow_tf32 = old


class _Block(torch.nn.Module):
    def __init__(self, c, k, fused):
        super().__init__()
        nn = pdl.nn
        self.conv1, self.bn1 = nn.Conv2d(c, c, k, 1, k // 2, bias=False), nn.BatchNorm2d(c)
        self.conv2, self.bn2 = nn.Conv2d(c, c, 1, bias=False), nn.BatchNorm2d(c)
        self.fused = fused

    def forward(self, x):
        if self.fused:
            y, skip = self.conv1(x, bn_stats=True, skip=True)
        else:
            y, skip = self.conv1(x), x
        y = self.bn1(y, relu=True)
        return self.bn2(self.conv2(y), relu=True, residual=skip)


@pytest.mark.parametrize("c,k,hw,n", [(64, 3, 14, 4), (128, 1, 9, 3), (256, 3, 7, 5), (64, 3, 30, 2)])
def test_skip_gradient_joins_in_the_dgrad_epilogue(c, k, hw, n, monkeypatch):
    torch.manual_seed(3)
    dev = torch.device("cuda", 0)
    a = _Block(c, k, True).to(dev)
    b = copy.deepcopy(a)
    b.fused = False
    pre = pdl.nn.Conv2d(c, c, 1, bias=False).to(dev)  # makes the block input a non-leaf activation with its own gradient
    pre_b = copy.deepcopy(pre)
    x = torch.randn(n, c, hw, hw, device=dev).contiguous(memory_format=torch.channels_last)
    for blk, p in ((a, pre), (b, pre_b)):
        out = blk(p(x))
        out.square().mean().backward()
    for (name, pa), (_, pb) in zip(list(a.named_parameters()) + list(pre.named_parameters()), list(b.named_parameters()) + list(pre_b.named_parameters())):
        ref = pb.grad.abs().max().item() + 1e-12
        err = (pa.grad - pb.grad).abs().max().item()
        assert err <= 2e-3 * ref, "%s: %.3e vs magnitude %.3e" % (name, err, ref)
    # and with the fused addend switched off the same module still works (plain in-place add)
    monkeypatch.setenv("PDL_FUSE_SKIP_ADD", "0")
    a.zero_grad()
    pre.zero_grad()
    a(pre(x)).square().mean().backward()
    ref = pre_b.weight.grad.abs().max().item()
    assert (pre.weight.grad - pre_b.weight.grad).abs().max().item() <= 2e-3 * ref
