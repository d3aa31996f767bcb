"""Dispatch to the tcgen05/TMA implicit-GEMM convolution kernels (csrc/cuda/tc_conv.cu, tc_wgrad.cu).

Each function returns True when a tensor-core kernel handled the call and False when the shape
is outside what those kernels cover (the caller then uses the CUDA-core kernel).  TF32 multiplies
with fp32 accumulation -- the precision the reference's convolutions run at by default -- so the
path is skipped when the user disables TF32 convs (``torch.backends.cudnn.allow_tf32 = False``)
or sets ``PDL_DISABLE_TC=1``.
"""
from __future__ import annotations

import ctypes
import os

import torch

from .. import _lib

_err = {}


def enabled():
    return torch.backends.cudnn.allow_tf32 and os.environ.get("PDL_DISABLE_TC", "0") != "1"


def _p(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def _stream(t):
    return ctypes.c_void_p(torch.cuda.current_stream(t.device).cuda_stream)


def error_flag(device):
    """Device int set by a kernel whose pipeline wait timed out (checked by tests / check_errors)."""
    f = _err.get(device.index)
    if f is None:
        f = _err[device.index] = torch.zeros(1, dtype=torch.int32, device=device)
    return f


def check_errors(device):
    f = _err.get(device.index)
    if f is not None and int(f.item()) != 0:
        f.zero_()
        raise RuntimeError("a tensor-core convolution kernel timed out waiting on its TMA/MMA pipeline")


def _ok(rc, what):
    if rc != 0:
        raise RuntimeError("%s failed (%d): %s" % (what, rc, _lib.cuda().pdl_last_error().decode()))


def pack_weights(w, want_fwd=True, want_bwd=False):
    """w[Cout][Cin][KH][KW] -> ([tap][Cout][Cin], [tap][Cin][Cout]) K-major operand copies."""
    Cout, Cin, KH, KW = w.shape
    fwd = torch.empty((KH * KW, Cout, Cin), dtype=w.dtype, device=w.device) if want_fwd else None
    bwd = torch.empty((KH * KW, Cin, Cout), dtype=w.dtype, device=w.device) if want_bwd else None
    _ok(_lib.cuda().pdl_tc_pack_weights(_p(w), _p(fwd), _p(bwd), Cout, Cin, KH * KW, _stream(w)), "weight packing")
    return fwd, bwd


def conv2d_fwd(x, w, b, y, stride, pad, relu=False):
    if not enabled():
        return False
    N, Cin, H, W = x.shape
    Cout, _, KH, KW = w.shape
    L = _lib.cuda()
    if not L.pdl_tc_conv_supported(H, W, Cin, Cout, KH, KW, stride, pad, 0):
        return False
    packed, _ = pack_weights(w, True, False)
    _ok(L.pdl_tc_conv2d_fwd_strided(_p(x), _p(packed), _p(b), _p(y), N, H, W, Cin, Cout, KH, KW, stride, pad, int(relu),
                                    _p(error_flag(x.device)), _stream(x)), "tensor-core conv forward")
    return True


def conv2d_fwd_relu_pool(x, w, b, pool_stride):
    """Fused conv + bias + ReLU + 2x2 max-pool.  Returns (pooled NHWC-as-NCHW tensor, argmax uint8) or None."""
    if not enabled():
        return None
    N, Cin, H, W = x.shape
    Cout, _, KH, KW = w.shape
    L = _lib.cuda()
    if not L.pdl_tc_conv_supported(H, W, Cin, Cout, KH, KW, 1, 0, 1):
        return None
    P, Q = H - KH + 1, W - KW + 1
    PH, PW = (P - 2) // pool_stride + 1, (Q - 2) // pool_stride + 1
    packed, _ = pack_weights(w, True, False)
    y = torch.empty((N, Cout, PH, PW), dtype=x.dtype, device=x.device, memory_format=torch.channels_last)
    idx = torch.empty((N, PH, PW, Cout), dtype=torch.uint8, device=x.device)
    _ok(L.pdl_tc_conv2d_fwd(_p(x), _p(packed), _p(b), _p(y), _p(idx), N, H, W, Cin, Cout, KH, KW, 0, 1, pool_stride,
                            _p(error_flag(x.device)), _stream(x)), "tensor-core conv+relu+pool forward")
    return y, idx


def conv2d_dgrad(dy, w, dx, stride, pad):
    if not enabled() or stride not in (1, 2):
        return False
    N, Cin, H, W = dx.shape
    Cout, _, KH, KW = w.shape
    if Cout % 32 or not (Cin in (32, 64) or Cin % 128 == 0) or W > 128 * stride:
        return False
    _, packed = pack_weights(w, False, True)
    _ok(_lib.cuda().pdl_tc_conv2d_dgrad_strided(_p(dy), _p(packed), _p(dx), N, H, W, Cin, Cout, KH, KW, stride, pad,
                                                _p(error_flag(dx.device)), _stream(dx)), "tensor-core conv dgrad")
    return True


def conv2d_wgrad(x, dy, dw, db, stride, pad):
    if not enabled():
        return False
    N, Cin, H, W = x.shape
    Cout, _, KH, KW = dw.shape
    L = _lib.cuda()
    if not L.pdl_tc_wgrad_supported(N, H, W, Cin, Cout, KH, KW, stride, pad):
        return False
    from .functional import scratch

    sc = scratch(L.pdl_tc_wgrad_scratch_bytes_strided(N, H, W, Cin, Cout, KH, KW, stride, pad), x.device, "tc_wgrad")
    db_part = None
    if db is not None:
        db_part = dy.sum(dim=(0, 2, 3)).reshape(1, Cout).contiguous()  # eager path only; the engine fuses this into pool-backward
    _ok(L.pdl_tc_conv2d_wgrad_strided(_p(x), _p(dy), _p(dw), _p(db), _p(db_part), 1 if db is not None else 0, _p(sc), N, H, W, Cin, Cout,
                                      KH, KW, stride, pad, _p(error_flag(x.device)), _stream(x)), "tensor-core conv wgrad")
    return True
