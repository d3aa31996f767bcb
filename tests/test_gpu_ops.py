"""Kernel parity on one B200: every native op against a plain PyTorch fp32 reference."""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

import pytorch_distributed_learning_b200 as pdl  # noqa: E402
from pytorch_distributed_learning_b200 import _lib  # noqa: E402
from pytorch_distributed_learning_b200.ops import functional as NF  # noqa: E402


@pytest.fixture(autouse=True)
def _strict_fp32():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def _dev():
    return torch.device("cuda", 0)


CONV_CASES = [
    # N, Cin, H, Cout, k, stride, pad  (ConvNet layers first, then ResNet-style shapes)
    (5, 1, 28, 32, 5, 1, 1), (5, 32, 13, 64, 3, 1, 0), (5, 64, 10, 128, 3, 1, 0),
    (3, 3, 32, 64, 7, 2, 3), (2, 64, 8, 128, 3, 2, 1), (2, 64, 8, 128, 1, 2, 0), (4, 16, 9, 24, 3, 1, 1),
]


@pytest.mark.parametrize("case", CONV_CASES)
def test_conv2d_fwd_bwd(case):
    N, Cin, H, Cout, k, s, p = case
    torch.manual_seed(0)
    x = torch.randn(N, Cin, H, H, device=_dev(), requires_grad=True)
    w = torch.randn(Cout, Cin, k, k, device=_dev(), requires_grad=True) * 0.1
    w.retain_grad()
    b = torch.randn(Cout, device=_dev(), requires_grad=True)
    xr, wr, br = (t.detach().clone().requires_grad_(True) for t in (x, w, b))
    y = NF.conv2d(x, w, b, s, p)
    yr = F.conv2d(xr, wr, br, stride=s, padding=p)
    torch.testing.assert_close(y, yr, rtol=1e-4, atol=1e-4)
    dy = torch.randn_like(yr)
    y.backward(dy)
    yr.backward(dy)
    torch.testing.assert_close(x.grad, xr.grad, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(w.grad, wr.grad, rtol=1e-4, atol=2e-4)
    torch.testing.assert_close(b.grad, br.grad, rtol=1e-4, atol=2e-4)


@pytest.mark.parametrize("k,s,p,H", [(2, 2, 0, 26), (2, 1, 0, 11), (2, 2, 0, 8), (3, 2, 1, 16), (3, 2, 1, 7)])
def test_maxpool_fwd_bwd(k, s, p, H):
    torch.manual_seed(1)
    x = torch.randn(4, 24, H, H, device=_dev(), requires_grad=True)
    xr = x.detach().clone().requires_grad_(True)
    y = NF.max_pool2d(x, k, s, p)
    yr = F.max_pool2d(xr, k, s, p)
    assert torch.equal(y, yr)
    dy = torch.randn_like(yr)
    y.backward(dy)
    yr.backward(dy)
    torch.testing.assert_close(x.grad, xr.grad, rtol=1e-6, atol=1e-6)


def test_maxpool_ties_follow_first_max_rule():
    x = torch.zeros(1, 4, 6, 6, device=_dev(), requires_grad=True)  # all ties (post-ReLU zeros)
    xr = x.detach().clone().requires_grad_(True)
    dy = torch.randn(1, 4, 5, 5, device=_dev())
    NF.max_pool2d(x, 2, 1, 0).backward(dy)
    F.max_pool2d(xr, 2, 1, 0).backward(dy)
    torch.testing.assert_close(x.grad, xr.grad)


def test_relu_linear_cross_entropy():
    torch.manual_seed(2)
    x = torch.randn(37, 2048, device=_dev(), requires_grad=True)
    w = (torch.randn(10, 2048, device=_dev()) * 0.02).requires_grad_(True)
    b = torch.randn(10, device=_dev(), requires_grad=True)
    t = torch.randint(0, 10, (37,), device=_dev())
    xr, wr, br = (v.detach().clone().requires_grad_(True) for v in (x, w, b))
    loss = NF.cross_entropy(NF.linear(NF.relu(x), w, b), t)
    lossr = F.cross_entropy(F.linear(F.relu(xr), wr, br), t)
    torch.testing.assert_close(loss, lossr, rtol=1e-5, atol=1e-5)
    (loss * 3.0).backward()
    (lossr * 3.0).backward()
    torch.testing.assert_close(x.grad, xr.grad, rtol=1e-4, atol=1e-6)
    torch.testing.assert_close(w.grad, wr.grad, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(b.grad, br.grad, rtol=1e-4, atol=1e-5)


def test_global_avgpool():
    x = torch.randn(3, 32, 5, 5, device=_dev(), requires_grad=True)
    xr = x.detach().clone().requires_grad_(True)
    y, yr = NF.global_avg_pool(x), F.adaptive_avg_pool2d(xr, 1)
    torch.testing.assert_close(y, yr)
    dy = torch.randn_like(yr)
    y.backward(dy)
    yr.backward(dy)
    torch.testing.assert_close(x.grad, xr.grad)


def test_convnet_modules_match_torch():
    torch.manual_seed(3)
    ours = pdl.models.ConvNet().to(_dev())

    class Ref(torch.nn.Module):  # launch_dist.py:9-41 written from its layer list
        def __init__(self):
            super().__init__()
            self.conv1, self.conv2, self.conv3 = torch.nn.Conv2d(1, 32, 5, 1, 1), torch.nn.Conv2d(32, 64, 3), torch.nn.Conv2d(64, 128, 3)
            self.fc1 = torch.nn.Linear(2048, 10)

        def forward(self, x):
            x = F.max_pool2d(torch.relu(self.conv1(x)), 2, 2)
            x = F.max_pool2d(torch.relu(self.conv2(x)), 2, 1)
            x = F.max_pool2d(torch.relu(self.conv3(x)), 2, 2)
            return self.fc1(x.view(-1, 2048))

    ref = Ref().to(_dev())
    ref.load_state_dict(ours.state_dict())
    x = torch.randn(9, 1, 28, 28, device=_dev())
    t = torch.randint(0, 10, (9,), device=_dev())
    lo = pdl.nn.CrossEntropyLoss()(ours(x), t)
    lr = torch.nn.CrossEntropyLoss()(ref(x), t)
    torch.testing.assert_close(lo, lr, rtol=1e-5, atol=1e-5)
    lo.backward()
    lr.backward()
    for (n, a), (_, b) in zip(ours.named_parameters(), ref.named_parameters()):
        torch.testing.assert_close(a.grad, b.grad, rtol=1e-3, atol=1e-5, msg=lambda m, n=n: "%s: %s" % (n, m))


@pytest.mark.parametrize("kw", [dict(), dict(momentum=0.9, weight_decay=1e-4, nesterov=True), dict(momentum=0.5, dampening=0.1)])
def test_multi_tensor_sgd_matches_torch(kw):
    torch.manual_seed(4)
    a = [torch.randn(s, device=_dev(), requires_grad=True) for s in [(5, 3), (4097,), (128, 64, 3, 3), (1,)]]
    b = [t.detach().clone().requires_grad_(True) for t in a]
    oa, ob = pdl.optim.SGD(a, lr=0.05, **kw), torch.optim.SGD(b, lr=0.05, **kw)
    for _ in range(3):
        for ta, tb in zip(a, b):
            g = torch.randn_like(ta)
            ta.grad, tb.grad = g.clone(), g.clone()
        oa.step()
        ob.step()
    for ta, tb in zip(a, b):
        torch.testing.assert_close(ta, tb, rtol=1e-6, atol=1e-6)


# ---------------------------------------------------------------------------- allreduce kernels
def _emulated_allreduce(algo, world, n, sgd, steps=2, blocks=3):
    """All `world` ranks as one cooperative launch on one GPU (see B200_PROFILING: ranks that wait
    on each other must not be separate launches)."""
    L = _lib.cuda()
    dev = _dev()
    n4 = (n + 3) // 4
    g = torch.Generator(device="cpu").manual_seed(5)
    ins = [torch.zeros(n4 * 4, device=dev) for _ in range(world)]
    outs = [torch.full((n4 * 4,), float("nan"), device=dev) for _ in range(world)]
    rings = [torch.zeros(n4 * 4 * (2 * world if algo == 4 else 1), device=dev) for _ in range(world)]  # push: the bucket's inbox
    pads = [torch.zeros(L.pdl_signal_pad_bytes() // 4, dtype=torch.int32, device=dev) for _ in range(world)]
    states = [torch.zeros(L.pdl_comm_state_bytes(), dtype=torch.uint8, device=dev) for _ in range(world)]
    p0 = torch.randn(n4 * 4, generator=g)
    params = [p0.clone().to(dev) for _ in range(world)]
    moms = [torch.zeros(n4 * 4, device=dev) for _ in range(world)]
    hyper = torch.zeros(8)
    hyper[0], hyper[1], hyper[2], hyper[3] = 0.1, 0.9, 0.0, 1e-3
    hyper.view(torch.int32)[4] = 1
    hyper.view(torch.int32)[5] = 1
    hypers = [hyper.clone().to(dev) for _ in range(world)]
    arr = lambda ts: (ctypes.c_void_p * len(ts))(*[t.data_ptr() for t in ts])  # noqa: E731
    ref_p, ref_m = p0.clone().double(), torch.zeros(n4 * 4).double()
    for step in range(steps):
        grads = [torch.randn(n4 * 4, generator=g) for _ in range(world)]
        for t, gr in zip(ins, grads):
            t.copy_(gr)
        rc = L.pdl_allreduce_emulated(algo, world, n4, arr(ins), arr(outs), arr(rings), arr(pads), arr(params) if sgd else None,
                                      arr(moms) if sgd else None, arr(hypers) if sgd else None, arr(states), blocks,
                                      ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, L.pdl_last_error()
        torch.cuda.synchronize()
        mean = torch.stack(grads).double().mean(0)
        for v in range(world):
            torch.testing.assert_close(outs[v].cpu().double(), mean, rtol=1e-5, atol=1e-6)
            assert torch.equal(outs[v], outs[0]), "replicas must receive bit-identical averages"
            assert int(states[v].view(torch.int32)[L.pdl_comm_state_error_offset() // 4].item()) == 0, "peer wait timed out"
        if sgd:
            gg = mean + 1e-3 * ref_p
            ref_m = gg.clone() if step == 0 else 0.9 * ref_m + gg
            ref_p = ref_p - 0.1 * (gg + 0.9 * ref_m)
            for v in range(world):
                torch.testing.assert_close(params[v].cpu().double(), ref_p, rtol=1e-4, atol=1e-5)
                assert torch.equal(params[v], params[0])


@pytest.mark.parametrize("algo", [0, 1, 2, 4])
@pytest.mark.parametrize("world", [2, 4, 8])
@pytest.mark.parametrize("n", [113674, 1000, 7])
def test_allreduce_variants_emulated(algo, world, n):
    _emulated_allreduce(algo, world, n, sgd=False)


@pytest.mark.parametrize("algo", [0, 1, 2, 4])
def test_allreduce_fused_sgd_emulated(algo):
    _emulated_allreduce(algo, 4, 113674, sgd=True, steps=3)


def test_allreduce_soak_emulated():
    _emulated_allreduce(0, 8, 4096, sgd=False, steps=200, blocks=4)
    _emulated_allreduce(2, 8, 4096, sgd=False, steps=100, blocks=4)
    _emulated_allreduce(4, 8, 4096, sgd=False, steps=201, blocks=4)  # push: odd count, so both inbox parities end up reused


@pytest.mark.parametrize("shape", [(8, 64, 8, 8), (4, 128, 4, 4), (16, 256, 2, 2), (3, 64, 16, 16), (64, 512, 1, 1)])
def test_batchnorm_train_fwd_bwd_matches_torch(shape):
    torch.manual_seed(6)
    N, C, H, W = shape
    x = (torch.randn(shape, device=_dev()) * 2 + 0.5).requires_grad_(True)
    xr = x.detach().clone().requires_grad_(True)
    ours, ref = pdl.nn.BatchNorm2d(C).to(_dev()), torch.nn.BatchNorm2d(C).to(_dev())
    with torch.no_grad():
        ours.weight.uniform_(0.5, 1.5)
        ours.bias.uniform_(-0.5, 0.5)
    ref.load_state_dict(ours.state_dict())
    for _ in range(2):  # twice: running statistics must keep tracking torch's
        y, yr = ours(x), ref(xr)
    torch.testing.assert_close(y, yr, rtol=1e-4, atol=1e-4)
    dy = torch.randn_like(yr)
    y.backward(dy)
    yr.backward(dy)
    torch.testing.assert_close(x.grad, xr.grad, rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(ours.weight.grad, ref.weight.grad, rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(ours.bias.grad, ref.bias.grad, rtol=1e-3, atol=1e-3)
    torch.testing.assert_close(ours.running_mean, ref.running_mean, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(ours.running_var, ref.running_var, rtol=1e-4, atol=1e-5)
    assert int(ours.num_batches_tracked) == int(ref.num_batches_tracked) == 2
    ours.eval(), ref.eval()
    with torch.no_grad():
        torch.testing.assert_close(ours(x), ref(xr), rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("relu,with_res", [(True, False), (True, True), (False, True)])
@pytest.mark.parametrize("shape", [(8, 64, 14, 14), (5, 96, 7, 9), (2, 256, 40, 40)])
def test_batchnorm_fused_relu_residual(shape, relu, with_res):
    """bn(x, relu=, residual=) against the unfused torch composition, forward, all gradients and eval mode."""
    torch.manual_seed(7)
    N, C, H, W = shape
    x = (torch.randn(shape, device=_dev()) * 1.5 - 0.3).requires_grad_(True)
    r = torch.randn(shape, device=_dev()).requires_grad_(True) if with_res else None
    xr = x.detach().clone().requires_grad_(True)
    rr = r.detach().clone().requires_grad_(True) if with_res else None
    ours, ref = pdl.nn.BatchNorm2d(C).to(_dev()), torch.nn.BatchNorm2d(C).to(_dev())
    with torch.no_grad():
        ours.weight.uniform_(0.5, 1.5)
        ours.bias.uniform_(-0.5, 0.5)
    ref.load_state_dict(ours.state_dict())

    def ref_form(inp, res):
        out = ref(inp)
        if res is not None:
            out = out + res
        return torch.relu(out) if relu else out

    y, yr = ours(x, relu=relu, residual=r), ref_form(xr, rr)
    torch.testing.assert_close(y, yr, rtol=1e-4, atol=1e-4)
    # keep the test away from the ReLU kink: elements whose pre-activation is ~0 may legitimately flip
    safe = (yr.detach().abs() > 1e-4) if relu else torch.ones_like(yr, dtype=torch.bool)
    dy = torch.randn_like(yr) * safe
    y.backward(dy)
    yr.backward(dy)
    torch.testing.assert_close(x.grad, xr.grad, rtol=1e-3, atol=2e-4)
    if with_res:
        torch.testing.assert_close(r.grad * safe, rr.grad * safe, rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(ours.weight.grad, ref.weight.grad, rtol=1e-3, atol=2e-3)
    torch.testing.assert_close(ours.bias.grad, ref.bias.grad, rtol=1e-3, atol=2e-3)
    torch.testing.assert_close(ours.running_var, ref.running_var, rtol=1e-4, atol=1e-5)
    ours.eval(), ref.eval()
    with torch.no_grad():
        torch.testing.assert_close(ours(x, relu=relu, residual=r), ref_form(xr, rr), rtol=1e-4, atol=1e-4)


def test_maxpool_vector_path_matches_torch():
    x = torch.randn(3, 64, 30, 30, device=_dev(), requires_grad=True)
    xr = x.detach().clone().requires_grad_(True)
    y, yr = pdl.nn.MaxPool2d(3, 2, 1)(x), torch.nn.functional.max_pool2d(xr, 3, 2, 1)
    torch.testing.assert_close(y, yr)
    dy = torch.randn_like(yr)
    y.backward(dy)
    yr.backward(dy)
    torch.testing.assert_close(x.grad, xr.grad)


def test_add_relu():
    a = torch.randn(4, 64, 8, 8, device=_dev(), requires_grad=True)
    b = torch.randn(4, 64, 8, 8, device=_dev(), requires_grad=True)
    ar, br = a.detach().clone().requires_grad_(True), b.detach().clone().requires_grad_(True)
    y, yr = NF.add_relu(a, b), torch.relu(ar + br)
    torch.testing.assert_close(y, yr)
    dy = torch.randn_like(yr)
    y.backward(dy)
    yr.backward(dy)
    torch.testing.assert_close(a.grad, ar.grad)
    torch.testing.assert_close(b.grad, br.grad)


# (N, Cin, H, Cout, k, stride, pad): whole-image tiles, row-strip tiles with a ragged last strip, several images per
# tile, several column tiles, stride 2, and the small-Cin stem
BN_STATS_CASES = [(4, 64, 56, 64, 3, 1, 1), (3, 64, 14, 256, 1, 1, 0), (5, 128, 7, 128, 3, 1, 1), (2, 256, 28, 512, 1, 2, 0),
                  (3, 128, 30, 128, 3, 1, 1), (6, 3, 64, 64, 7, 2, 3)]


@pytest.mark.parametrize("case", BN_STATS_CASES)
def test_conv_epilogue_batchnorm_statistics(case):
    """conv(x, bn_stats=True) -> bn: the statistics come from the convolution's epilogue; same result as bn's own pass."""
    import copy

    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True  # the tensor-core path is what carries the fused statistics
    try:
        N, Cin, H, Cout, k, stride, pad = case
        torch.manual_seed(12)
        conv = pdl.nn.Conv2d(Cin, Cout, k, stride, pad, bias=False).to(_dev())
        bn_a = pdl.nn.BatchNorm2d(Cout).to(_dev())
        with torch.no_grad():
            bn_a.weight.uniform_(0.5, 1.5)
            bn_a.bias.uniform_(-0.5, 0.5)
        bn_b = copy.deepcopy(bn_a)
        x = (torch.randn(N, Cin, H, H, device=_dev()) + 0.3).requires_grad_(Cin > 4)
        ya = conv(x, bn_stats=True)
        partial = getattr(ya, "_pdl_bn_partial", None)
        assert partial is not None and partial.shape[1:] == (Cout, 3)
        P = (H + 2 * pad - k) // stride + 1
        assert torch.equal(partial[:, :, 0].sum(0), torch.full((Cout,), float(N * P * P), device=_dev()))
        out_a = bn_a(ya, relu=True)
        yb = conv(x)
        assert not hasattr(yb, "_pdl_bn_partial")
        torch.testing.assert_close(ya, yb, rtol=0, atol=0)  # same kernel, same tile order
        out_b = bn_b(yb, relu=True)
        torch.testing.assert_close(out_a, out_b, rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(bn_a.running_mean, bn_b.running_mean, rtol=1e-4, atol=1e-6)
        torch.testing.assert_close(bn_a.running_var, bn_b.running_var, rtol=1e-4, atol=1e-6)
        assert int(bn_a.num_batches_tracked) == 1
        out_a.square().sum().backward()
        assert conv.weight.grad is not None and torch.isfinite(conv.weight.grad).all()
    finally:
        torch.backends.cudnn.allow_tf32 = old
