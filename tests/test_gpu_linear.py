"""Classifier-head Linear (long K, small batch) against strict-fp32 torch: forward, dgrad, wgrad."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from pytorch_distributed_learning_b200.ops import functional as NF  # noqa: E402


@pytest.fixture(autouse=True)
def _strict_fp32():
    old = torch.backends.cuda.matmul.allow_tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cuda.matmul.allow_tf32 = old


@pytest.mark.parametrize("M,N,K,bias", [(64, 1000, 2048, True), (37, 1000, 2048, True), (128, 96, 512, False), (5, 64, 1024, True), (300, 130, 640, True)])
def test_linear_rows_kernel_matches_torch(M, N, K, bias):
    torch.manual_seed(0)
    dev = torch.device("cuda", 0)
    x = torch.randn(M, K, device=dev, requires_grad=True)
    w = (torch.randn(N, K, device=dev) * 0.02).requires_grad_(True)
    b = torch.randn(N, device=dev, requires_grad=True) if bias else None
    xr, wr = x.detach().clone().requires_grad_(True), w.detach().clone().requires_grad_(True)
    br = b.detach().clone().requires_grad_(True) if bias else None
    y, yr = NF.linear(x, w, b), F.linear(xr, wr, br)
    torch.testing.assert_close(y, yr, rtol=1e-4, atol=1e-4)
    dy = torch.randn_like(yr)
    y.backward(dy)
    yr.backward(dy)
    torch.testing.assert_close(x.grad, xr.grad, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(w.grad, wr.grad, rtol=1e-4, atol=1e-4)
    if bias:
        torch.testing.assert_close(b.grad, br.grad, rtol=1e-4, atol=1e-4)
