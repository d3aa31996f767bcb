"""tcgen05 / TMA convolution kernels against the strict-fp32 CUDA-core oracle and PyTorch."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from pytorch_distributed_learning_b200 import _lib  # noqa: E402
from pytorch_distributed_learning_b200.ops import functional as NF  # noqa: E402
from pytorch_distributed_learning_b200.ops import tc_conv  # noqa: E402

DEV = torch.device("cuda", 0)
CL = torch.channels_last


@pytest.fixture(autouse=True)
def _tf32_convs():
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True  # torch's default, and what enables the tensor-core path here
    yield
    torch.backends.cudnn.allow_tf32 = old


def _tf32_close(a, b, scale_ref=None):
    # TF32 multiplies: ~1e-3 relative to the magnitude of the accumulated products
    ref = (scale_ref if scale_ref is not None else b).abs().max().item() + 1e-6
    err = (a - b).abs().max().item()
    assert err <= 4e-3 * ref, "max abs err %.3e vs ref magnitude %.3e" % (err, ref)


# (N, Cin, H, Cout, k, pad): ConvNet conv2 / conv3 and a few other stride-1 shapes
FWD_CASES = [(5, 32, 13, 64, 3, 0), (5, 64, 10, 128, 3, 0), (100, 32, 13, 64, 3, 0), (7, 64, 10, 128, 3, 0),
             (3, 64, 8, 64, 3, 1), (2, 32, 20, 32, 3, 1), (4, 128, 6, 128, 1, 0)]


@pytest.mark.parametrize("case", FWD_CASES)
def test_tc_forward_matches_fp32(case):
    N, Cin, H, Cout, k, pad = case
    torch.manual_seed(0)
    x = torch.randn(N, Cin, H, H, device=DEV).contiguous(memory_format=CL)
    w = torch.randn(Cout, Cin, k, k, device=DEV) * 0.1
    b = torch.randn(Cout, device=DEV)
    P = H + 2 * pad - k + 1
    y = torch.empty((N, Cout, P, P), device=DEV).contiguous(memory_format=CL)
    assert tc_conv.conv2d_fwd(x, w, b, y, 1, pad)
    tc_conv.check_errors(DEV)
    torch.backends.cudnn.allow_tf32 = False
    ref = F.conv2d(x, w, b, padding=pad)
    torch.backends.cudnn.allow_tf32 = True
    _tf32_close(y, ref)


@pytest.mark.parametrize("case,ps", [((5, 32, 13, 64, 3, 0), 1), ((6, 64, 10, 128, 3, 0), 2), ((7, 64, 10, 128, 3, 0), 2), ((100, 32, 13, 64, 3, 0), 1)])
def test_tc_forward_fused_relu_pool(case, ps):
    N, Cin, H, Cout, k, pad = case
    torch.manual_seed(1)
    x = torch.randn(N, Cin, H, H, device=DEV).contiguous(memory_format=CL)
    w = torch.randn(Cout, Cin, k, k, device=DEV) * 0.1
    b = torch.randn(Cout, device=DEV)
    out = tc_conv.conv2d_fwd_relu_pool(x, w, b, ps)
    assert out is not None
    y, idx = out
    tc_conv.check_errors(DEV)
    torch.backends.cudnn.allow_tf32 = False
    z = F.relu(F.conv2d(x, w, b))
    torch.backends.cudnn.allow_tf32 = True
    ref = F.max_pool2d(z, 2, ps)
    _tf32_close(y, ref)
    # the argmax must point at an element that attains the pooled value of OUR conv output
    assert int(idx.max()) <= 3


DGRAD_CASES = [(5, 32, 13, 64, 3, 0), (5, 64, 10, 128, 3, 0), (100, 64, 10, 128, 3, 0), (3, 64, 8, 64, 3, 1)]


@pytest.mark.parametrize("case", DGRAD_CASES)
def test_tc_dgrad_matches_fp32(case):
    N, Cin, H, Cout, k, pad = case
    torch.manual_seed(2)
    P = H + 2 * pad - k + 1
    dy = torch.randn(N, Cout, P, P, device=DEV).contiguous(memory_format=CL)
    w = torch.randn(Cout, Cin, k, k, device=DEV) * 0.1
    dx = torch.empty((N, Cin, H, H), device=DEV).contiguous(memory_format=CL)
    assert tc_conv.conv2d_dgrad(dy, w, dx, 1, pad)
    tc_conv.check_errors(DEV)
    torch.backends.cudnn.allow_tf32 = False
    ref = torch.nn.grad.conv2d_input((N, Cin, H, H), w, dy, padding=pad)
    torch.backends.cudnn.allow_tf32 = True
    _tf32_close(dx, ref)


WGRAD_CASES = [(5, 32, 13, 64, 3, 0), (5, 64, 10, 128, 3, 0), (100, 32, 13, 64, 3, 0), (100, 64, 10, 128, 3, 0), (3, 64, 8, 64, 3, 1)]


@pytest.mark.parametrize("case", WGRAD_CASES)
def test_tc_wgrad_matches_fp32(case):
    N, Cin, H, Cout, k, pad = case
    torch.manual_seed(3)
    P = H + 2 * pad - k + 1
    x = torch.randn(N, Cin, H, H, device=DEV).contiguous(memory_format=CL)
    dy = torch.randn(N, Cout, P, P, device=DEV).contiguous(memory_format=CL)
    dw = torch.empty(Cout, Cin, k, k, device=DEV)
    db = torch.empty(Cout, device=DEV)
    assert tc_conv.conv2d_wgrad(x, dy, dw, db, 1, pad)
    tc_conv.check_errors(DEV)
    torch.backends.cudnn.allow_tf32 = False
    ref = torch.nn.grad.conv2d_weight(x, (Cout, Cin, k, k), dy, padding=pad)
    torch.backends.cudnn.allow_tf32 = True
    _tf32_close(dw, ref)
    torch.testing.assert_close(db, dy.sum(dim=(0, 2, 3)), rtol=1e-4, atol=1e-3)


def test_tc_path_is_used_by_autograd_function():
    torch.manual_seed(4)
    x = torch.randn(4, 32, 13, 13, device=DEV, requires_grad=True)
    w = (torch.randn(64, 32, 3, 3, device=DEV) * 0.1).requires_grad_(True)
    b = torch.randn(64, device=DEV, requires_grad=True)
    xr, wr, br = (t.detach().clone().requires_grad_(True) for t in (x, w, b))
    assert tc_conv.enabled()
    y = NF.conv2d(x, w, b, 1, 0)
    yr = F.conv2d(xr, wr, br)
    dy = torch.randn_like(yr)
    y.backward(dy)
    yr.backward(dy)
    tc_conv.check_errors(DEV)
    _tf32_close(y, yr)
    _tf32_close(x.grad, xr.grad)
    _tf32_close(w.grad, wr.grad)
