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
             (3, 64, 8, 64, 3, 1), (2, 32, 20, 32, 3, 1), (4, 128, 6, 128, 1, 0),
             # ResNet-style: wide channels (column tiles), 1x1, large spatial extents (row-strip tiles)
             (2, 256, 4, 256, 3, 1), (2, 512, 2, 512, 3, 1), (2, 64, 56, 256, 1, 0), (2, 128, 28, 128, 3, 1), (3, 1024, 14, 256, 1, 0)]


@pytest.mark.parametrize("case", FWD_CASES)
def test_tc_forward_matches_fp32(case):
    N, Cin, H, Cout, k, pad = case
    torch.manual_seed(0)
    x = torch.randn(N, Cin, H, H, device=DEV).contiguous(memory_format=CL)
    w = torch.randn(Cout, Cin, k, k, device=DEV) * 0.1
    b = torch.randn(Cout, device=DEV)
    P = H + 2 * pad - k + 1
    y = torch.empty((N, Cout, P, P), device=DEV).contiguous(memory_format=CL)
    assert tc_conv.conv2d_fwd(x, w, b, y, 1, pad)[0]
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
    # the arg-max byte (row << 1 | col inside the 2x2 window) must select an element that attains the pooled value: gather the
    # strict-fp32 conv output at the selected position and compare it with the pooled reference (TF32 tolerance, which also
    # covers windows whose two best candidates are closer than the rounding error).  The backward pass routes gradients by it.
    assert int(idx.max()) <= 3
    win = z.unfold(2, 2, ps).unfold(3, 2, ps).reshape(N, Cout, ref.shape[2], ref.shape[3], 4)  # [n][c][ph][pw][r*2+s]
    sel = torch.gather(win, 4, idx.permute(0, 3, 1, 2).long().unsqueeze(-1)).squeeze(-1)
    _tf32_close(sel, ref)
    live = ref > 0.05  # pooled value clearly positive: the selected element must be that same (positive) maximum
    assert bool((sel[live] > 0).all())


DGRAD_CASES = [(5, 32, 13, 64, 3, 0), (5, 64, 10, 128, 3, 0), (100, 64, 10, 128, 3, 0), (3, 64, 8, 64, 3, 1),
               (2, 256, 4, 256, 3, 1), (2, 256, 56, 64, 1, 0), (2, 128, 28, 128, 3, 1), (3, 256, 14, 1024, 1, 0)]


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


WGRAD_CASES = [(5, 32, 13, 64, 3, 0), (5, 64, 10, 128, 3, 0), (100, 32, 13, 64, 3, 0), (100, 64, 10, 128, 3, 0), (3, 64, 8, 64, 3, 1),
               (2, 256, 4, 256, 3, 1), (2, 64, 56, 256, 1, 0), (2, 128, 28, 128, 3, 1), (3, 1024, 14, 256, 1, 0), (4, 512, 7, 512, 3, 1)]


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


@pytest.mark.parametrize("case", [(2, 64, 56, 128, 3, 1), (3, 256, 14, 512, 1, 0), (2, 128, 28, 256, 3, 1), (5, 64, 8, 128, 3, 1), (4, 512, 7, 1024, 1, 0)])
def test_tc_forward_stride2(case):
    N, Cin, H, Cout, k, pad = case
    torch.manual_seed(5)
    x = torch.randn(N, Cin, H, H, device=DEV).contiguous(memory_format=CL)
    w = torch.randn(Cout, Cin, k, k, device=DEV) * 0.1
    P = (H + 2 * pad - k) // 2 + 1
    y = torch.empty((N, Cout, P, P), device=DEV).contiguous(memory_format=CL)
    assert tc_conv.conv2d_fwd(x, w, None, y, 2, pad)[0]
    tc_conv.check_errors(DEV)
    torch.backends.cudnn.allow_tf32 = False
    ref = F.conv2d(x, w, None, stride=2, padding=pad)
    torch.backends.cudnn.allow_tf32 = True
    _tf32_close(y, ref)


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


@pytest.mark.parametrize("variant", ["default", "single-stream", "fused-fc-pool2"])
@pytest.mark.parametrize("batch", [100, 7])
@pytest.mark.parametrize("opt_kw", [dict(lr=1e-2), dict(lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)])
def test_graphed_trainstep_tensor_core_path(batch, opt_kw, variant, monkeypatch):
    """The flagship path: whole step in one CUDA graph with the tcgen05 conv kernels (TF32), against fp32 torch.
    Variants: the default forked graph, everything on one stream, and the opt-in fused kernels."""
    import copy

    if variant == "single-stream":
        monkeypatch.setenv("PDL_CONVNET_FORK", "0")
    elif variant == "fused-fc-pool2":
        monkeypatch.setenv("PDL_CONVNET_FUSE_FC", "1")
        monkeypatch.setenv("PDL_CONVNET_FUSE_POOL2", "1")

    import pytorch_distributed_learning_b200 as pdl
    from test_gpu_engine import _reference_run

    torch.manual_seed(0)
    model = pdl.models.ConvNet().cuda()
    state = copy.deepcopy(model.state_dict())
    opt = pdl.optim.SGD(model.parameters(), **opt_kw)
    step = pdl.TrainStep(model, pdl.nn.CrossEntropyLoss(), opt)
    assert step.graphed
    g = torch.Generator().manual_seed(1)
    batches = [(torch.randn(batch, 1, 28, 28, generator=g).pin_memory(), torch.randint(0, 10, (batch,), generator=g).pin_memory()) for _ in range(5)]
    losses = [step(x, y).item() for x, y in batches]
    assert step.engine.use_tc and step.kernels_per_step() <= 16
    tc_conv.check_errors(DEV)
    torch.backends.cudnn.allow_tf32 = False
    ref, ref_losses = _reference_run(state, batches, opt_kw)
    torch.backends.cudnn.allow_tf32 = True
    ref_tf32, _ = _reference_run(state, batches, opt_kw)  # the reference's DEFAULT precision (cuDNN TF32 convs)
    torch.testing.assert_close(torch.tensor(losses), torch.tensor(ref_losses), rtol=3e-3, atol=3e-3)
    # parity criterion: be as close to strict fp32 as torch's own TF32 convolutions are (see profiles/tf32_parity.md)
    for (n, a), (_, b), (_, c) in zip(model.named_parameters(), ref.named_parameters(), ref_tf32.named_parameters()):
        delta_ref = (b - state[n].cuda()).norm().item()
        err_ours, err_torch = (a - b).norm().item(), (c - b).norm().item()
        assert err_ours <= 2.5 * err_torch + 1e-2 * delta_ref + 1e-7, \
            "%s: |ours - fp32| = %.3e, |torch TF32 - fp32| = %.3e, update norm %.3e" % (n, err_ours, err_torch, delta_ref)


# (N, Cin, H (pooled extent), Cout, k, pad, pool stride)
POOLBWD_CASES = [(5, 64, 10, 128, 3, 0, 1), (100, 64, 10, 128, 3, 0, 1), (7, 32, 6, 64, 3, 1, 2), (9, 64, 5, 64, 3, 1, 1), (3, 128, 8, 128, 1, 0, 2)]


@pytest.mark.parametrize("case", POOLBWD_CASES)
def test_tc_dgrad_with_fused_pool_backward(case):
    """conv dgrad whose epilogue also runs the 2x2 max-pool backward + ReLU mask + bias partials of the layer below,
    against the same thing done by torch ops on strict-fp32 data."""
    N, Cin, H, Cout, k, pad, ps = case
    torch.manual_seed(11)
    L = _lib.cuda()
    ZH = (H - 1) * ps + 2
    z = torch.randn(N, Cin, ZH, ZH, device=DEV)                     # pre-activation of the layer below
    a = torch.relu(z)
    pooled, flat_idx = F.max_pool2d(a, 2, ps, return_indices=True)    # [N, Cin, H, H]
    # arg-max code 0..3 inside each window, NHWC like the kernels store it
    ph = torch.arange(H, device=DEV).view(1, 1, H, 1) * ps
    pw = torch.arange(H, device=DEV).view(1, 1, 1, H) * ps
    code = ((flat_idx // ZH - ph) * 2 + (flat_idx % ZH - pw)).to(torch.uint8)
    idx_nhwc = code.permute(0, 2, 3, 1).contiguous()
    pooled_nhwc = pooled.contiguous(memory_format=CL)
    P = H + 2 * pad - k + 1
    dy = torch.randn(N, Cout, P, P, device=DEV).contiguous(memory_format=CL)
    w = torch.randn(Cout, Cin, k, k, device=DEV) * 0.1
    _, packed = tc_conv.pack_weights(w, False, True)
    tiles = L.pdl_tc_dgrad_poolbwd_tiles(N, H, H)
    assert tiles > 0
    dz = torch.full((N, ZH, ZH, Cin), float("nan"), device=DEV)
    dbpart = torch.full((tiles, Cin), float("nan"), device=DEV)
    rc = L.pdl_tc_conv2d_dgrad_poolbwd(NF._p(dy), NF._p(packed), NF._p(idx_nhwc), NF._p(pooled_nhwc), NF._p(dz), NF._p(dbpart), N, H, H, Cin, Cout,
                                       k, k, pad, ps, NF._p(tc_conv.error_flag(DEV)), NF._stream(dy))
    assert rc == 0, L.pdl_last_error()
    tc_conv.check_errors(DEV)
    # oracle: strict-fp32 dgrad, then autograd through relu + max_pool
    torch.backends.cudnn.allow_tf32 = False
    da = torch.nn.grad.conv2d_input((N, Cin, H, H), w, dy, stride=1, padding=pad)
    torch.backends.cudnn.allow_tf32 = True
    zr = z.clone().requires_grad_(True)
    F.max_pool2d(torch.relu(zr), 2, ps).backward(da)
    ref = zr.grad.permute(0, 2, 3, 1)
    _tf32_close(dz, ref, scale_ref=da)
    torch.testing.assert_close(dbpart.sum(0), dz.sum((0, 1, 2)), rtol=2e-3, atol=2e-3 * dz.abs().sum((0, 1, 2)).max().item() / 10 + 1e-4)


def test_convnet_eager_fused_blocks_match_torch():
    """The reference's unchanged forward (launch_dist.py:35-41) on this package's layers takes the fused conv+ReLU+pool kernels
    (``ops.functional.conv_relu_pool``) when TF32 convolutions are allowed; loss and every gradient must match torch's
    strict-fp32 result within TF32 tolerance, and the bias gradients (pool-backward partials) exactly as closely."""
    import copy

    import pytorch_distributed_learning_b200 as pdl

    torch.manual_seed(3)
    ours = pdl.models.ConvNet().to(DEV)

    class Ref(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.conv1, self.conv2, self.conv3 = torch.nn.Conv2d(1, 32, 5, 1, 1), torch.nn.Conv2d(32, 64, 3), torch.nn.Conv2d(64, 128, 3)
            self.fc1 = torch.nn.Linear(2048, 10)

        def forward(self, x):
            x = F.max_pool2d(torch.relu(self.conv1(x)), 2, 2)
            x = F.max_pool2d(torch.relu(self.conv2(x)), 2, 1)
            x = F.max_pool2d(torch.relu(self.conv3(x)), 2, 2)
            return self.fc1(x.reshape(-1, 2048))

    ref = Ref().to(DEV)
    ref.load_state_dict(copy.deepcopy(ours.state_dict()))
    x, y = torch.randn(20, 1, 28, 28, device=DEV), torch.randint(0, 10, (20,), device=DEV)
    assert tc_conv.enabled()
    before = _lib.cuda().pdl_launch_count()
    lo = pdl.nn.CrossEntropyLoss()(ours(x), y)
    lo.backward()
    launched = _lib.cuda().pdl_launch_count() - before
    tc_conv.check_errors(DEV)
    torch.backends.cudnn.allow_tf32 = False
    lr = F.cross_entropy(ref(x), y)
    lr.backward()
    strict = [p.grad.clone() for p in ref.parameters()]
    torch.backends.cudnn.allow_tf32 = True  # torch's own TF32 convolutions: the yardstick for "TF32 tolerance" through 3 layers
    ref.zero_grad()
    F.cross_entropy(ref(x), y).backward()
    assert abs(lo.item() - lr.item()) < 2e-3 * max(1.0, abs(lr.item()))
    for (n, a), b, t in zip(ours.named_parameters(), strict, [p.grad for p in ref.parameters()]):
        mag = b.abs().max().item()
        ours_err, torch_err = (a.grad - b).abs().max().item(), (t - b).abs().max().item()
        assert ours_err <= max(3.0 * torch_err, 4e-3 * mag), "%s: ours %.3e vs torch-TF32 %.3e (|g| max %.3e)" % (n, ours_err, torch_err, mag)
    assert launched <= 30, "the fused path should need ~25 launches, saw %d" % launched


def test_cta_pair_variant_matches_single_cta():
    """The opt-in CTA-pair kernel (cluster of 2, weight tile TMA-multicast to both CTAs, multicast commit on the stage barriers)
    must give bit-identical results to the single-CTA kernel: same tiles, same MMAs, only the way the weight tile arrives in
    shared memory differs.  Runs in a child process because the choice is read from PDL_TC_CLUSTER once per process."""
    import os
    import subprocess
    import sys

    code = r'''
import torch, sys
sys.path.insert(0, ".")
from pytorch_distributed_learning_b200.ops import tc_conv
torch.manual_seed(0)
res = []
for (N, Cin, H, Cout, k, pad) in [(8, 128, 28, 128, 3, 1), (4, 64, 56, 256, 1, 0), (6, 256, 14, 512, 1, 0)]:
    x = torch.randn(N, Cin, H, H, device="cuda").contiguous(memory_format=torch.channels_last)
    w = torch.randn(Cout, Cin, k, k, device="cuda") * 0.1
    y = torch.empty((N, Cout, H, H), device="cuda").contiguous(memory_format=torch.channels_last)
    handled, _, _ = tc_conv.conv2d_fwd(x, w, None, y, 1, pad, relu=False, want_bwd_pack=False, bn_stats=False)
    assert handled
    dx = torch.empty_like(x)
    assert tc_conv.conv2d_dgrad(y, w, dx, 1, pad)
    tc_conv.check_errors(x.device)
    res += [y.cpu(), dx.cpu()]
torch.save(res, sys.argv[1])
'''
    import tempfile

    outs = []
    for mode in ("0", "2"):
        with tempfile.NamedTemporaryFile(suffix=".pt") as f:
            env = dict(os.environ, PDL_TC_CLUSTER=mode)
            subprocess.run([sys.executable, "-c", code, f.name], check=True, env=env, cwd=os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
            outs.append(torch.load(f.name))
    for a, b in zip(*outs):
        assert torch.equal(a, b)
