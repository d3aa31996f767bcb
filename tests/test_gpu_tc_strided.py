"""Stride-2 tcgen05 convolutions (ResNet down-sampling layers): forward uses the TMA traversal stride, dgrad is four
parity-class launches, wgrad reads the input through a strided box.  Checked against strict-fp32 PyTorch."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from pytorch_distributed_learning_b200.ops import functional as NF  # noqa: E402
from pytorch_distributed_learning_b200.ops import tc_conv  # noqa: E402

DEV = torch.device("cuda", 0)
CL = torch.channels_last


@pytest.fixture(autouse=True)
def _tf32_convs():
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True
    yield
    torch.backends.cudnn.allow_tf32 = old


def _close(a, b):
    ref = b.abs().max().item() + 1e-6
    err = (a - b).abs().max().item()
    assert err <= 4e-3 * ref, "max abs err %.3e vs ref magnitude %.3e" % (err, ref)


# (N, Cin, H, Cout, k, pad) with stride 2: ResNet-18/50 down-sampling shapes, incl. odd extents
CASES = [(2, 64, 56, 128, 3, 1), (3, 256, 14, 512, 1, 0), (2, 128, 28, 256, 3, 1), (5, 64, 8, 128, 3, 1), (4, 512, 7, 1024, 1, 0),
         (3, 64, 9, 64, 3, 1), (2, 256, 56, 512, 1, 0)]


@pytest.mark.parametrize("case", CASES)
def test_dgrad_stride2(case):
    N, Cin, H, Cout, k, pad = case
    torch.manual_seed(7)
    P = (H + 2 * pad - k) // 2 + 1
    dy = torch.randn(N, Cout, P, P, device=DEV).contiguous(memory_format=CL)
    w = torch.randn(Cout, Cin, k, k, device=DEV) * 0.1
    dx = torch.full((N, Cin, H, H), float("nan"), device=DEV).contiguous(memory_format=CL)
    assert tc_conv.conv2d_dgrad(dy, w, dx, 2, pad)
    tc_conv.check_errors(DEV)
    torch.backends.cudnn.allow_tf32 = False
    ref = torch.nn.grad.conv2d_input((N, Cin, H, H), w, dy, stride=2, padding=pad)
    torch.backends.cudnn.allow_tf32 = True
    _close(dx, ref)


@pytest.mark.parametrize("case", CASES)
def test_wgrad_stride2(case):
    N, Cin, H, Cout, k, pad = case
    torch.manual_seed(8)
    P = (H + 2 * pad - k) // 2 + 1
    x = torch.randn(N, Cin, H, H, device=DEV).contiguous(memory_format=CL)
    dy = torch.randn(N, Cout, P, P, device=DEV).contiguous(memory_format=CL)
    dw = torch.empty(Cout, Cin, k, k, device=DEV)
    assert tc_conv.conv2d_wgrad(x, dy, dw, None, 2, pad)
    tc_conv.check_errors(DEV)
    torch.backends.cudnn.allow_tf32 = False
    ref = torch.nn.grad.conv2d_weight(x, (Cout, Cin, k, k), dy, stride=2, padding=pad)
    torch.backends.cudnn.allow_tf32 = True
    _close(dw, ref)


def test_autograd_conv_stride2_uses_tensor_cores():
    torch.manual_seed(9)
    x = torch.randn(4, 64, 16, 16, device=DEV, requires_grad=True)
    w = (torch.randn(128, 64, 3, 3, device=DEV) * 0.1).requires_grad_(True)
    xr, wr = x.detach().clone().requires_grad_(True), w.detach().clone().requires_grad_(True)
    y, yr = NF.conv2d(x, w, None, 2, 1), F.conv2d(xr, wr, None, stride=2, padding=1)
    dy = torch.randn_like(yr)
    y.backward(dy)
    yr.backward(dy)
    tc_conv.check_errors(DEV)
    _close(y, yr)
    _close(x.grad, xr.grad)
    _close(w.grad, wr.grad)


# (N, Cin, H, W, Cout, k, stride, pad): small-Cin first layers; the first is ResNet's 7x7/2 stem at ImageNet size
STEMS = [(2, 3, 224, 224, 64, 7, 2, 3), (3, 3, 32, 32, 64, 7, 2, 3), (2, 1, 28, 28, 64, 5, 1, 1), (2, 3, 32, 32, 64, 3, 1, 1),
         (1, 4, 37, 53, 128, 5, 2, 2)]


@pytest.mark.parametrize("case", STEMS)
@pytest.mark.parametrize("channels_last", [False, True])
def test_stem_conv_fwd_wgrad(case, channels_last):
    """Cin <= 4 convolutions on the tensor cores through the padded NHWC4 virtual layout (overlapping TMA windows)."""
    N, Cin, H, W, Cout, k, stride, pad = case
    torch.manual_seed(9)
    x = torch.randn(N, Cin, H, W, device=DEV)
    if channels_last:
        x = x.contiguous(memory_format=CL)
    w = (torch.randn(Cout, Cin, k, k, device=DEV) * 0.1).requires_grad_(True)
    assert tc_conv.stem_supported(x, w, stride, pad)
    y = NF.conv2d(x, w, None, stride, pad)
    tc_conv.check_errors(DEV)
    dy = torch.randn_like(y)
    y.backward(dy)
    tc_conv.check_errors(DEV)
    torch.backends.cudnn.allow_tf32 = False
    wr = w.detach().clone().requires_grad_(True)
    yr = F.conv2d(x, wr, None, stride, pad)
    yr.backward(dy)
    torch.backends.cudnn.allow_tf32 = True
    _close(y, yr)
    _close(w.grad, wr.grad)
