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
    """Word that a tensor-core kernel sets when one of its (bounded) pipeline waits timed out.  It lives in PINNED HOST memory,
    which the device can address directly: the failing kernel writes it itself, and ``check_errors`` is a plain host load --
    cheap enough for the step path of the graphed engines, no synchronisation, no copy."""
    f = _err.get(device.index)
    if f is None:
        f = _err[device.index] = torch.zeros(1, dtype=torch.int32).pin_memory()
    return f


def check_errors(device):
    """Raise if a tensor-core convolution kernel gave up on its TMA/MMA pipeline since the last check (its output is garbage)."""
    f = _err.get(device.index)
    if f is not None and int(f[0]) != 0:
        f.zero_()
        raise RuntimeError("a tensor-core convolution kernel timed out waiting on its TMA/MMA pipeline")


def _ok(rc, what):
    if rc != 0:
        raise RuntimeError("%s failed (%d): %s" % (what, rc, _lib.cuda().pdl_last_error().decode()))


_prepack_cache = {}   # id(model) -> persistent buffers + pointer table of its convolution weights
_prepack_live = {}    # weight data_ptr -> (fwd, bwd) while a `prepacked(model)` block is open


class prepacked:
    """``with prepacked(model): forward; backward`` -- packs EVERY tensor-core-eligible convolution weight of the model
    with one kernel launch up front; ``pack_weights`` then hands out those copies instead of launching per layer
    (a per-layer pack is a ~5 us latency-bound kernel: 52 of them in a ResNet-50 step).  The weights must not change
    inside the block (they do not: the optimizer runs after backward, and the fused allreduce+SGD kernels only touch
    the raw parameters, never the packed copies).  Used by the graph engine around each captured iteration."""

    def __init__(self, model):
        self.model = model

    def _entry(self):
        key = id(self.model)
        ent = _prepack_cache.get(key)
        ws = [m.weight for m in self.model.modules()
              if isinstance(m, torch.nn.Conv2d) and m.weight.is_cuda and m.weight.dtype == torch.float32 and m.weight.dim() == 4
              and m.groups == 1 and m.weight.shape[1] % 32 == 0 and (m.weight.shape[0] in (32, 64) or m.weight.shape[0] % 128 == 0)
              and m.weight.shape[2] * m.weight.shape[3] <= _lib.cuda().pdl_tc_pack_max_taps() and m.weight.is_contiguous()]
        sig = tuple(w.data_ptr() for w in ws)
        if ent is not None and ent["sig"] == sig:
            return ent
        if not ws:
            ent = {"sig": sig, "num": 0}
        else:
            L = _lib.cuda()
            dev = ws[0].device
            rows, prefix, packs, total = [], [0], {}, 0
            for w in ws:
                Cout, Cin, KH, KW = w.shape
                fwd = torch.empty((KH * KW, Cout, Cin), dtype=w.dtype, device=dev)
                bwd = torch.empty((KH * KW, Cin, Cout), dtype=w.dtype, device=dev)
                packs[w.data_ptr()] = (fwd, bwd)
                rows.append([w.data_ptr(), fwd.data_ptr(), bwd.data_ptr(), Cout, Cin, KH * KW])
                total += (Cout // 32) * (Cin // 32)  # one CTA per tile of 32 x 32 channels
                prefix.append(total)
            ent = {"sig": sig, "num": len(ws), "packs": packs, "total": total,
                   "host": (torch.tensor(rows, dtype=torch.int64).pin_memory(), torch.tensor(prefix, dtype=torch.int32).pin_memory())}
            # pinned staging + async copies: legal even if the first use happens under stream capture
            ent["table"], ent["prefix"] = ent["host"][0].to(dev, non_blocking=True), ent["host"][1].to(dev, non_blocking=True)
        _prepack_cache[key] = ent
        return ent

    def __enter__(self):
        if not enabled() or os.environ.get("PDL_PREPACK", "1") == "0":
            return self
        ent = self._entry()
        if ent["num"]:
            dev = ent["table"].device
            _ok(_lib.cuda().pdl_tc_pack_weights_multi(_p(ent["table"]), _p(ent["prefix"]), ent["num"], ent["total"],
                                                      ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)), "weight packing (all layers)")
            _prepack_live.update(ent["packs"])
        return self

    def __exit__(self, *exc):
        _prepack_live.clear()
        return False


def pack_weights(w, want_fwd=True, want_bwd=False):
    """w[Cout][Cin][KH][KW] -> ([tap][Cout][Cin], [tap][Cin][Cout]) K-major operand copies."""
    live = _prepack_live.get(w.data_ptr())
    if live is not None and live[0].shape[1] == w.shape[0] and live[0].shape[2] == w.shape[1]:
        return (live[0] if want_fwd else None), (live[1] if want_bwd else None)
    Cout, Cin, KH, KW = w.shape
    fwd = torch.empty((KH * KW, Cout, Cin), dtype=w.dtype, device=w.device) if want_fwd else None
    bwd = torch.empty((KH * KW, Cin, Cout), dtype=w.dtype, device=w.device) if want_bwd else None
    _ok(_lib.cuda().pdl_tc_pack_weights(_p(w), _p(fwd), _p(bwd), Cout, Cin, KH * KW, _stream(w)), "weight packing")
    return fwd, bwd


def dgrad_supported(Cin, Cout, W, stride):
    return stride in (1, 2) and Cout % 32 == 0 and (Cin in (32, 64) or Cin % 128 == 0) and W <= 128 * stride


def conv2d_fwd(x, w, b, y, stride, pad, relu=False, want_bwd_pack=False, bn_stats=False):
    """Returns (handled, packed_bwd, bn_partial).  With ``want_bwd_pack`` the dgrad operand copy of the weights is produced
    by the same packing launch and handed back for the backward pass (None when dgrad will not use the tensor cores).  With
    ``bn_stats`` (bias-free, no ReLU) the kernel's epilogue also reduces the BatchNorm statistics of y: bn_partial is
    [splits][Cout][3] (count, mean, M2), or None when that is not possible."""
    if not enabled():
        return False, None, None
    N, Cin, H, W = x.shape
    Cout, _, KH, KW = w.shape
    L = _lib.cuda()
    if not L.pdl_tc_conv_supported(H, W, Cin, Cout, KH, KW, stride, pad, 0):
        return False, None, None
    packed, packed_bwd = pack_weights(w, True, want_bwd_pack and dgrad_supported(Cin, Cout, W, stride))
    splits = L.pdl_tc_conv_bn_splits(N, H, W, Cin, Cout, KH, KW, stride, pad) if (bn_stats and b is None and not relu and fused_bn_stats()) else 0
    if splits > 0:
        partial = torch.empty((splits, Cout, 3), dtype=torch.float32, device=x.device)
        _ok(L.pdl_tc_conv2d_fwd_bnstats(_p(x), _p(packed), _p(y), _p(partial), N, H, W, Cin, Cout, KH, KW, stride, pad,
                                        _p(error_flag(x.device)), _stream(x)), "tensor-core conv forward + BatchNorm statistics")
        return True, packed_bwd, partial
    _ok(L.pdl_tc_conv2d_fwd_strided(_p(x), _p(packed), _p(b), _p(y), N, H, W, Cin, Cout, KH, KW, stride, pad, int(relu),
                                    _p(error_flag(x.device)), _stream(x)), "tensor-core conv forward")
    return True, packed_bwd, None


def fused_bn_stats():
    return os.environ.get("PDL_FUSE_BN_STATS", "1") != "0"


def stem_supported(x, w, stride, pad):
    """Small-Cin first layer (Cin <= 4, e.g. ResNet's 7x7/2 conv1): forward and wgrad both on the tensor cores."""
    if not enabled():
        return False
    N, Cin, H, W = x.shape
    Cout, _, KH, KW = w.shape
    L = _lib.cuda()
    return bool(Cin <= 4 and L.pdl_tc_stem_supported(H, W, Cin, Cout, KH, KW, stride, pad)
                and L.pdl_tc_stem_wgrad_supported(N, H, W, Cin, Cout, KH, KW, stride, pad))


def stem_fwd(x, w, b, y, stride, pad, bn_stats=False):
    """x in ANY layout (strides are passed down) -> y (NHWC); returns (the padded NHWC4 copy the wgrad reuses, BatchNorm
    statistics partials of y or None)."""
    N, Cin, H, W = x.shape
    Cout, _, KH, KW = w.shape
    L = _lib.cuda()
    xp = torch.empty(L.pdl_tc_stem_pad_floats(N, H, W, KH, KW, stride, pad), dtype=x.dtype, device=x.device)
    sn, sc, sh, sw = x.stride()
    _ok(L.pdl_tc_stem_prepare(_p(x), sn, sc, sh, sw, _p(xp), N, Cin, H, W, KH, KW, stride, pad, _stream(x)), "stem input padding")
    packed = torch.empty((KH, Cout, 32), dtype=w.dtype, device=w.device)
    _ok(L.pdl_tc_stem_pack_weights(_p(w), _p(packed), Cout, Cin, KH, KW, _stream(x)), "stem weight packing")
    splits = L.pdl_tc_stem_bn_splits(N, H, W, Cin, Cout, KH, KW, stride, pad) if (bn_stats and b is None and fused_bn_stats()) else 0
    if splits > 0:
        partial = torch.empty((splits, Cout, 3), dtype=torch.float32, device=x.device)
        _ok(L.pdl_tc_stem_fwd_bnstats(_p(xp), _p(packed), _p(y), _p(partial), N, H, W, Cout, KH, KW, stride, pad, _p(error_flag(x.device)),
                                      _stream(x)), "tensor-core stem conv forward + BatchNorm statistics")
        return xp, partial
    _ok(L.pdl_tc_stem_fwd(_p(xp), _p(packed), _p(b), _p(y), N, H, W, Cout, KH, KW, stride, pad, 0, _p(error_flag(x.device)), _stream(x)),
        "tensor-core stem conv forward")
    return xp, None


def stem_wgrad(xp, dy, dw, x_shape, stride, pad):
    N, Cin, H, W = x_shape
    Cout, _, KH, KW = dw.shape
    L = _lib.cuda()
    from .functional import scratch

    sc = scratch(L.pdl_tc_stem_wgrad_scratch_bytes(N, H, W, Cin, Cout, KH, KW, stride, pad), dy.device, "tc_wgrad")
    _ok(L.pdl_tc_stem_wgrad(_p(xp), _p(dy), _p(dw), _p(sc), N, H, W, Cin, Cout, KH, KW, stride, pad, _p(error_flag(dy.device)), _stream(dy)),
        "tensor-core stem conv wgrad")


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


def conv2d_dgrad(dy, w, dx, stride, pad, packed=None, add=None):
    """dx = dgrad(dy, w) [+ add].  Returns False when the tensor cores do not cover the shape, True when dx is complete,
    and "no-add" when dx was written WITHOUT the addend (the caller adds it)."""
    N, Cin, H, W = dx.shape
    Cout, _, KH, KW = w.shape
    if not enabled() or not dgrad_supported(Cin, Cout, W, stride):
        return False
    if packed is None:
        _, packed = pack_weights(w, False, True)
    if add is not None and stride == 1 and add.shape == dx.shape and add.is_contiguous(memory_format=torch.channels_last) \
            and os.environ.get("PDL_FUSE_SKIP_ADD", "1") != "0":
        rc = _lib.cuda().pdl_tc_conv2d_dgrad_add(_p(dy), _p(packed), _p(add), _p(dx), N, H, W, Cin, Cout, KH, KW, pad,
                                                 _p(error_flag(dx.device)), _stream(dx))
        if rc == 0:
            return True
        if rc != -5:
            _ok(rc, "tensor-core conv dgrad + skip gradient")
    _ok(_lib.cuda().pdl_tc_conv2d_dgrad_strided(_p(dy), _p(packed), _p(dx), N, H, W, Cin, Cout, KH, KW, stride, pad,
                                                _p(error_flag(dx.device)), _stream(dx)), "tensor-core conv dgrad")
    return True if add is None else "no-add"


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
    db_part, parts = None, 0
    if db is not None:  # bias gradient: per-split column sums of dy, folded in by the weight-gradient kernel's second pass
        M = dy.shape[0] * dy.shape[2] * dy.shape[3]
        parts = L.pdl_colsum_splits(M, Cout)
        db_part = torch.empty((parts, Cout), dtype=dy.dtype, device=dy.device)
        _ok(L.pdl_colsum_partial(_p(dy), _p(db_part), M, Cout, _stream(dy)), "bias gradient partial sums")
    _ok(L.pdl_tc_conv2d_wgrad_strided(_p(x), _p(dy), _p(dw), _p(db), _p(db_part), parts, _p(sc), N, H, W, Cin, Cout,
                                      KH, KW, stride, pad, _p(error_flag(x.device)), _stream(x)), "tensor-core conv wgrad")
    return True
