"""The allocation-free wgrad entry point, driven through the C ABI exactly as an embedding host would.

Hosts that may not allocate when they need a weight gradient (processes whose device memory is fully
owned by someone else's pool, loops that must not risk the device synchronisation an allocator call can
cause) call ``pdl_conv2d_wgrad_simt_noalloc``: its split-K partial sums live in a fixed 8 MiB array inside
the library.  It must give the same numbers as ``pdl_conv2d_wgrad_simt`` (caller-provided scratch), also on
a second call that finds the previous call's leftovers in the array, and must refuse shapes that do
not fit.  The comparison is against the scratch-taking twin, not against torch: the twin itself is
checked against a strict-fp32 torch oracle in tests/test_gpu_ops.py.  This file sorts last on purpose: a ``-x`` run has already executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

from pytorch_distributed_learning_b200 import _lib  # noqa: E402

P, I = ctypes.c_void_p, ctypes.c_int


def _lib_with_sigs():
    L = _lib.cuda()
    L.pdl_conv2d_wgrad_simt_noalloc.argtypes = [P, P, P, P, I, I, I, I, I, I, I, I, I, P]
    L.pdl_conv2d_wgrad_simt_noalloc.restype = I
    L.pdl_conv2d_wgrad_noalloc_max_bytes.argtypes = []
    L.pdl_conv2d_wgrad_noalloc_max_bytes.restype = ctypes.c_longlong
    return L


def _run_both(L, N, Cin, H, Cout, k, stride, pad, seed):
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cpu").manual_seed(seed)
    Pq = (H + 2 * pad - k) // stride + 1
    x = torch.randn(N, H, H, Cin, generator=g).to(dev)     # NHWC
    dy = torch.randn(N, Pq, Pq, Cout, generator=g).to(dev)  # NHWC
    stream = P(torch.cuda.current_stream().cuda_stream)
    ptr = lambda t: P(t.data_ptr())  # noqa: E731
    need = L.pdl_conv2d_wgrad_scratch_bytes(N, H, H, Cin, Cout, k, k, stride, pad)
    scratch = torch.empty(need, dtype=torch.uint8, device=dev)
    dw_a, db_a = torch.empty(Cout, Cin, k, k, device=dev), torch.empty(Cout, device=dev)
    dw_b, db_b = torch.full_like(dw_a, float("nan")), torch.full_like(db_a, float("nan"))
    assert L.pdl_conv2d_wgrad_simt(ptr(x), ptr(dy), ptr(dw_a), ptr(db_a), ptr(scratch), N, H, H, Cin, Cout, k, k, stride, pad, stream) == 0
    rc = L.pdl_conv2d_wgrad_simt_noalloc(ptr(x), ptr(dy), ptr(dw_b), ptr(db_b), N, H, H, Cin, Cout, k, k, stride, pad, stream)
    assert rc == 0, L.pdl_last_error()
    torch.cuda.synchronize()
    return (dw_a, db_a), (dw_b, db_b)


def test_noalloc_wgrad_matches_scratch_version_twice_in_a_row():
    L = _lib_with_sigs()
    # two consecutive calls with different shapes and data: the second finds the first one's partial sums in the array
    for seed, shape in enumerate([(6, 32, 13, 64, 3, 1, 0), (5, 1, 28, 32, 5, 1, 1)]):
        (dw_a, db_a), (dw_b, db_b) = _run_both(L, *shape, seed=seed)
        torch.testing.assert_close(dw_b, dw_a, rtol=1e-4, atol=2e-4)
        torch.testing.assert_close(db_b, db_a, rtol=1e-4, atol=2e-4)


def test_noalloc_wgrad_rejects_shapes_beyond_the_builtin_buffer():
    L = _lib_with_sigs()
    N, Cin, H, Cout, k = 4096, 64, 10, 128, 3
    assert L.pdl_conv2d_wgrad_scratch_bytes(N, H, H, Cin, Cout, k, k, 1, 0) > L.pdl_conv2d_wgrad_noalloc_max_bytes()
    one = torch.zeros(1, device="cuda")
    rc = L.pdl_conv2d_wgrad_simt_noalloc(P(one.data_ptr()), P(one.data_ptr()), P(one.data_ptr()), None, N, H, H, Cin, Cout, k, k, 1, 0,
                                         P(torch.cuda.current_stream().cuda_stream))
    assert rc != 0 and b"built-in buffer" in L.pdl_last_error()
