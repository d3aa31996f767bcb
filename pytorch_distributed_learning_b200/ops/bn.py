"""BatchNorm2d on the native kernels (csrc/cuda/bn.cu): training-mode batch statistics + running-stat
update, eval-mode affine, and the closed-form backward.  NHWC / channels-last fp32.

Used by ``nn.BatchNorm2d`` for the ResNets the reference trains (example_launch.py:26)."""
from __future__ import annotations

import ctypes

import torch

from .. import _lib
from .functional import _nhwc, _ok, _p, _stream, scratch

_CL = torch.channels_last


class _BatchNormTrain(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, bias, running_mean, running_var, momentum, eps):
        L = _lib.cuda()
        x = _nhwc(x)
        N, C, H, W = x.shape
        M = N * H * W
        y = torch.empty_like(x, memory_format=_CL)
        save_mean = torch.empty(C, dtype=x.dtype, device=x.device)
        save_invstd = torch.empty(C, dtype=x.dtype, device=x.device)
        sc = scratch(L.pdl_bn_scratch_floats(M, C) * 4, x.device, "bn")
        _ok(L.pdl_bn_fwd_train(_p(x), _p(weight), _p(bias), _p(running_mean), _p(running_var), float(momentum), float(eps), _p(y), _p(save_mean),
                               _p(save_invstd), _p(sc), M, C, 0, _stream(x)), "batch norm forward")
        ctx.save_for_backward(x, weight, save_mean, save_invstd)
        return y

    @staticmethod
    def backward(ctx, dy):
        L = _lib.cuda()
        x, weight, save_mean, save_invstd = ctx.saved_tensors
        dy = _nhwc(dy)
        N, C, H, W = x.shape
        M = N * H * W
        dx = torch.empty_like(x, memory_format=_CL)
        dgamma = torch.empty(C, dtype=x.dtype, device=x.device)
        dbeta = torch.empty(C, dtype=x.dtype, device=x.device)
        sc = scratch(L.pdl_bn_scratch_floats(M, C) * 4, x.device, "bn")
        _ok(L.pdl_bn_bwd(_p(x), _p(dy), _p(weight), _p(save_mean), _p(save_invstd), _p(dx), _p(dgamma), _p(dbeta), _p(sc), M, C, _stream(x)),
            "batch norm backward")
        return dx, dgamma, dbeta, None, None, None, None


def batch_norm(x, weight, bias, running_mean, running_var, training, momentum, eps):
    C = x.shape[1]
    if C % 4 != 0:
        return torch.nn.functional.batch_norm(x, running_mean, running_var, weight, bias, training, momentum, eps)
    if training:
        return _BatchNormTrain.apply(x, weight, bias, running_mean, running_var, momentum, eps)
    # eval mode: an affine map of the input; gradients (if any) flow through plain tensor ops
    if torch.is_grad_enabled() and (x.requires_grad or weight.requires_grad):
        return torch.nn.functional.batch_norm(x, running_mean, running_var, weight, bias, False, momentum, eps)
    L = _lib.cuda()
    x = _nhwc(x)
    N, _, H, W = x.shape
    y = torch.empty_like(x, memory_format=_CL)
    sc = scratch(L.pdl_bn_scratch_floats(N * H * W, C) * 4, x.device, "bn")
    _ok(L.pdl_bn_fwd_eval(_p(x), _p(weight), _p(bias), _p(running_mean), _p(running_var), float(eps), _p(y), _p(sc), N * H * W, C, 0, _stream(x)),
        "batch norm (eval)")
    return y
