"""Out-of-bounds write checks for the kernels, without compute-sanitizer.

``compute-sanitizer`` is closed on this GPU pool (``profiles/compute_sanitizer_closed.log``), so the memcheck tier of the
test plan ("Race/hang") is replaced by guard zones: every output a kernel writes is carved out of a larger
allocation whose margins hold a byte pattern; after the kernel has run the margins must be untouched and the result must
still match the reference (so the kernel really wrote where it was told to).  Shapes are the awkward ones: ragged last
tiles, odd batch sizes, extents that do not fill a TMA box, tails that are not a multiple of the vector width.
"""
import ctypes

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from pytorch_distributed_learning_b200 import _lib  # noqa: E402
from pytorch_distributed_learning_b200.ops import tc_conv  # noqa: E402
from pytorch_distributed_learning_b200.ops.functional import _p, _stream  # noqa: E402

DEV = "cuda"
PAD = 8192
CL = torch.channels_last


@pytest.fixture(autouse=True)
def _tf32_default():
    old = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = True
    yield
    torch.backends.cudnn.allow_tf32 = old


class Guarded:
    """A tensor of ``shape`` (logical NCHW stored NHWC when ``nhwc``) inside a canary-filled allocation."""

    def __init__(self, shape, dtype=torch.float32, nhwc=False):
        item = torch.empty((), dtype=dtype).element_size()
        n = 1
        for d in shape:
            n *= d
        self.buf = torch.full((n * item + 2 * PAD,), 0xA5, dtype=torch.uint8, device=DEV)
        body = self.buf[PAD:PAD + n * item].view(dtype)
        if nhwc:
            N, C, H, W = shape
            self.t = body.view(N, H, W, C).permute(0, 3, 1, 2)
        else:
            self.t = body.view(*shape)
        self.nbytes = n * item

    def check(self, what):
        torch.cuda.synchronize()
        lo, hi = self.buf[:PAD], self.buf[PAD + self.nbytes:]
        assert bool((lo == 0xA5).all()), "%s: wrote BEFORE its output buffer" % what
        assert bool((hi == 0xA5).all()), "%s: wrote PAST its output buffer" % what


def _close(a, b, tol=4e-3):
    ref = b.abs().max().item() + 1e-6
    assert (a - b).abs().max().item() <= tol * ref


@pytest.mark.parametrize("case", [(3, 32, 13, 64, 3, 0), (5, 64, 9, 128, 3, 1), (2, 128, 30, 128, 3, 1), (3, 256, 15, 512, 1, 0), (7, 64, 57, 64, 3, 1)])
def test_tensor_core_conv_outputs_stay_in_bounds(case):
    N, Cin, H, Cout, k, pad = case
    torch.manual_seed(0)
    P = H + 2 * pad - k + 1
    x = torch.randn(N, Cin, H, H, device=DEV).contiguous(memory_format=CL)
    w = torch.randn(Cout, Cin, k, k, device=DEV) * 0.1
    y = Guarded((N, Cout, P, P), nhwc=True)
    handled, _, _ = tc_conv.conv2d_fwd(x, w, None, y.t, 1, pad, relu=False, want_bwd_pack=False, bn_stats=False)
    assert handled
    y.check("tc forward")
    torch.backends.cudnn.allow_tf32 = False
    ref = F.conv2d(x, w, padding=pad)
    _close(y.t, ref)
    dy = torch.randn_like(ref).contiguous(memory_format=CL)
    torch.backends.cudnn.allow_tf32 = True
    dx = Guarded((N, Cin, H, H), nhwc=True)
    if tc_conv.conv2d_dgrad(dy, w, dx.t, 1, pad):  # shapes the tensor-core dgrad does not cover fall back elsewhere
        dx.check("tc dgrad")
        torch.backends.cudnn.allow_tf32 = False
        _close(dx.t, torch.nn.grad.conv2d_input(x.shape, w, dy, padding=pad))
        torch.backends.cudnn.allow_tf32 = True
    dw, db = Guarded((Cout, Cin, k, k)), Guarded((Cout,))
    if tc_conv.conv2d_wgrad(x, dy, dw.t, db.t, 1, pad):
        dw.check("tc wgrad")
        db.check("tc wgrad bias")
        torch.backends.cudnn.allow_tf32 = False
        _close(dw.t, torch.nn.grad.conv2d_weight(x, w.shape, dy, padding=pad))
        _close(db.t, dy.sum(dim=(0, 2, 3)), tol=1e-4)
    tc_conv.check_errors(torch.device(DEV))


@pytest.mark.parametrize("N", [1, 7, 100])
def test_fused_convnet_kernels_stay_in_bounds(N):
    """conv+ReLU+pool forward (tcgen05 and the conv1 CUDA-core kernel), pool backward with bias partials, conv1 wgrad."""
    L = _lib.cuda()
    torch.manual_seed(1)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    x = torch.randn(N, 28, 28, 1, device=DEV)
    w1, b1 = torch.randn(32, 1, 5, 5, device=DEV) * 0.2, torch.randn(32, device=DEV) * 0.1
    a1, i1 = Guarded((N, 13, 13, 32)), Guarded((N, 13, 13, 32), dtype=torch.uint8)
    assert L.pdl_conv1_fused_fwd(_p(x), _p(w1), _p(b1), _p(a1.t), _p(i1.t), N, st) == 0
    a1.check("conv1 fused forward")
    i1.check("conv1 fused forward argmax")
    ref1 = F.max_pool2d(torch.relu(F.conv2d(x.permute(0, 3, 1, 2), w1, b1, padding=1)), 2, 2).permute(0, 2, 3, 1)
    _close(a1.t, ref1)
    # conv2 with fused ReLU + overlapping 2x2/1 pool
    w2, b2 = torch.randn(64, 32, 3, 3, device=DEV) * 0.1, torch.randn(64, device=DEV) * 0.1
    packed, _ = tc_conv.pack_weights(w2, True, False)
    a2, i2 = Guarded((N, 10, 10, 64)), Guarded((N, 10, 10, 64), dtype=torch.uint8)
    err = tc_conv.error_flag(torch.device(DEV))
    a1c = a1.t.contiguous()
    assert L.pdl_tc_conv2d_fwd(_p(a1c), _p(packed), _p(b2), _p(a2.t), _p(i2.t), N, 13, 13, 32, 64, 3, 3, 0, 1, 1, _p(err), st) == 0
    a2.check("tc conv + relu + pool forward")
    i2.check("tc conv + relu + pool argmax")
    torch.backends.cudnn.allow_tf32 = False
    ref2 = F.max_pool2d(torch.relu(F.conv2d(a1c.permute(0, 3, 1, 2), w2, b2)), 2, 1).permute(0, 2, 3, 1)
    torch.backends.cudnn.allow_tf32 = True
    _close(a2.t, ref2)
    # pool backward (ReLU mask folded in) + bias-gradient partials
    da2 = torch.randn(N, 10, 10, 64, device=DEV)
    blocks = L.pdl_pool2_bwd_blocks(N, 11, 11, 64)
    dz2, dbp = Guarded((N, 11, 11, 64)), Guarded((blocks, 64))
    a2c, i2c = a2.t.contiguous(), i2.t.contiguous()
    assert L.pdl_pool2_bwd_dbias(_p(da2), _p(i2c), _p(a2c), _p(dz2.t), _p(dbp.t), N, 11, 11, 64, 1, st) == 0
    dz2.check("pool backward")
    dbp.check("pool backward bias partials")
    # (dz2 is TF32-rounded for the tensor cores, the partial sums are not: equal up to that rounding)
    torch.testing.assert_close(dbp.t.sum(0), dz2.t.sum(dim=(0, 1, 2)), rtol=2e-2, atol=2e-2 * max(1.0, N ** 0.5))
    # conv1 weight gradient with pool1 backward folded into the operand load
    da1 = torch.randn(N, 13, 13, 32, device=DEV)
    gw1, gb1, part = Guarded((32, 1, 5, 5)), Guarded((32,)), Guarded((L.pdl_bwd_blocks(N), 832))
    i1c = i1.t.contiguous()
    assert L.pdl_conv1_wgrad_fused(_p(x), _p(da1), _p(i1c), _p(a1c), _p(gw1.t), _p(gb1.t), _p(part.t), N, st) == 0
    for g, name in ((gw1, "conv1 wgrad"), (gb1, "conv1 bias grad"), (part, "conv1 wgrad partial rows")):
        g.check(name)
    tc_conv.check_errors(torch.device(DEV))


@pytest.mark.parametrize("M,C", [(37, 64), (1000, 256), (5, 2048)])
def test_batchnorm_and_reductions_stay_in_bounds(M, C):
    L = _lib.cuda()
    torch.manual_seed(2)
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    x, res = torch.randn(M, C, device=DEV), torch.randn(M, C, device=DEV)
    gamma, beta = torch.rand(C, device=DEV) + 0.5, torch.randn(C, device=DEV)
    rm, rv, nbt = torch.zeros(C, device=DEV), torch.ones(C, device=DEV), torch.zeros((), dtype=torch.int64, device=DEV)
    sc = torch.empty(int(L.pdl_bn_scratch_floats(M, C)), device=DEV)
    y, saved, mask = Guarded((M, C)), Guarded((4, C)), Guarded((M * C // 4,), dtype=torch.uint8)
    assert L.pdl_bn_fwd_train(_p(x), _p(res), _p(gamma), _p(beta), _p(rm), _p(rv), _p(nbt), ctypes.c_float(0.1), ctypes.c_float(1e-5), _p(y.t), _p(saved.t),
                              _p(mask.t), _p(sc), M, C, 1, st) == 0
    for g, name in ((y, "bn forward"), (saved, "bn saved statistics"), (mask, "bn relu mask")):
        g.check(name)
    ref = torch.relu(F.batch_norm(x, None, None, gamma, beta, True, 0.1, 1e-5) + res)
    torch.testing.assert_close(y.t, ref, rtol=1e-4, atol=1e-4)
    dy, dy2 = torch.randn(M, C, device=DEV), torch.randn(M, C, device=DEV)
    dx, dres, dg, dbt = Guarded((M, C)), Guarded((M, C)), Guarded((C,)), Guarded((C,))
    assert L.pdl_bn_bwd2(_p(x), _p(dy), _p(dy2), _p(mask.t.contiguous()), _p(gamma), _p(saved.t.contiguous()), _p(dx.t), _p(dres.t), _p(dg.t), _p(dbt.t),
                         _p(sc), M, C, 1, st) == 0
    for g, name in ((dx, "bn backward dx"), (dres, "bn backward dres"), (dg, "bn dgamma"), (dbt, "bn dbeta")):
        g.check(name)
    # the two-addend form equals the one-addend form on the summed gradient
    dx1, dres1, dg1, db1 = torch.empty(M, C, device=DEV), torch.empty(M, C, device=DEV), torch.empty(C, device=DEV), torch.empty(C, device=DEV)
    both = dy + dy2
    assert L.pdl_bn_bwd(_p(x), _p(both), _p(mask.t.contiguous()), _p(gamma), _p(saved.t.contiguous()), _p(dx1), _p(dres1), _p(dg1), _p(db1), _p(sc), M, C, 1,
                        st) == 0
    torch.testing.assert_close(dx.t, dx1, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(dg.t, dg1, rtol=1e-4, atol=1e-3)
    # column sums (bias-gradient partials)
    parts = L.pdl_colsum_splits(M, C)
    cs = Guarded((parts, C))
    assert L.pdl_colsum_partial(_p(x), _p(cs.t), M, C, st) == 0
    cs.check("column-sum partials")
    torch.testing.assert_close(cs.t.sum(0), x.sum(0), rtol=1e-4, atol=1e-3)


@pytest.mark.parametrize("algo", [0, 1, 2, 4])
@pytest.mark.parametrize("n", [7, 4099])
def test_allreduce_variants_stay_in_bounds(algo, n):
    """Every peer-memory allreduce variant, 4 virtual ranks as one cooperative launch: the averaged buckets, and for the
    push variant the inboxes, are guarded."""
    L = _lib.cuda()
    world, blocks = 4, 3
    n4 = (n + 3) // 4
    torch.manual_seed(3)
    ins = [torch.randn(n4 * 4, device=DEV) for _ in range(world)]
    outs = [Guarded((n4 * 4,)) for _ in range(world)]
    rings = [Guarded((n4 * 4 * (2 * world if algo == 4 else 1),)) for _ in range(world)]
    for r in rings:
        r.t.zero_()
    pads = [torch.zeros(L.pdl_signal_pad_bytes() // 4, dtype=torch.int32, device=DEV) for _ in range(world)]
    states = [torch.zeros(L.pdl_comm_state_bytes(), dtype=torch.uint8, device=DEV) for _ in range(world)]
    arr = lambda ts: (ctypes.c_void_p * len(ts))(*[t.data_ptr() for t in ts])  # noqa: E731
    for _ in range(3):
        rc = L.pdl_allreduce_emulated(algo, world, n4, arr(ins), arr([o.t for o in outs]), arr([r.t for r in rings]), arr(pads), None, None, None,
                                      arr(states), blocks, ctypes.c_void_p(torch.cuda.current_stream().cuda_stream))
        assert rc == 0, L.pdl_last_error()
    for v in range(world):
        outs[v].check("allreduce output (algo %d)" % algo)
        rings[v].check("allreduce staging / inbox (algo %d)" % algo)
        torch.testing.assert_close(outs[v].t, torch.stack(ins).mean(0), rtol=1e-5, atol=1e-6)
