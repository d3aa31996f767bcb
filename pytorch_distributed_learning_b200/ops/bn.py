"""BatchNorm2d on the native kernels (csrc/cuda/bn.cu) with the ResNet block's ReLU and residual add
fused in: ``y = act(bn(x) [+ residual])``.  Training-mode batch statistics + running-stat update,
eval-mode affine, and the closed-form backward (which also returns the skip connection's gradient).
NHWC / channels-last fp32.

Used by ``nn.BatchNorm2d`` for the ResNets the reference trains (example_launch.py:26)."""
from __future__ import annotations

import torch

from .. import _lib
from .functional import _nhwc, _ok, _p, _stream, scratch

_CL = torch.channels_last


class _BatchNormAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, weight, bias, running_mean, running_var, num_batches_tracked, momentum, eps, relu, partial=None):
        L = _lib.cuda()
        x = _nhwc(x)
        if residual is not None:
            residual = _nhwc(residual)
        N, C, H, W = x.shape
        M = N * H * W
        y = torch.empty_like(x, memory_format=_CL)
        saved = torch.empty((4, C), dtype=x.dtype, device=x.device)
        ctx.relu, ctx.has_res = bool(relu), residual is not None
        # backward recomputes the ReLU mask from x, unless a residual was added: then the forward kernel leaves 4 sign
        # bits per 4 outputs (1/16 of the tensor) so that the output itself need not be re-read
        mask = torch.empty(M * C // 4, dtype=torch.uint8, device=x.device) if (ctx.relu and ctx.has_res) else None
        if partial is not None:  # statistics already reduced by the producing convolution's epilogue
            _ok(L.pdl_bn_fwd_train_prestats(_p(x), _p(residual), _p(weight), _p(bias), _p(running_mean), _p(running_var), _p(num_batches_tracked),
                                            float(momentum), float(eps), _p(y), _p(saved), _p(mask), _p(partial), partial.shape[0], M, C,
                                            int(relu), _stream(x)), "batch norm forward (statistics from the convolution)")
        else:
            sc = scratch(L.pdl_bn_scratch_floats(M, C) * 4, x.device, "bn")
            _ok(L.pdl_bn_fwd_train(_p(x), _p(residual), _p(weight), _p(bias), _p(running_mean), _p(running_var), _p(num_batches_tracked),
                                   float(momentum), float(eps), _p(y), _p(saved), _p(mask), _p(sc), M, C, int(relu), _stream(x)), "batch norm forward")
        ctx.save_for_backward(x, weight, saved, mask)
        return y

    @staticmethod
    def backward(ctx, dy):
        L = _lib.cuda()
        x, weight, saved, mask = ctx.saved_tensors
        dy = _nhwc(dy)
        N, C, H, W = x.shape
        M = N * H * W
        dx = torch.empty_like(x, memory_format=_CL)
        dres = torch.empty_like(x, memory_format=_CL) if ctx.has_res else None
        dgamma = torch.empty(C, dtype=x.dtype, device=x.device)
        dbeta = torch.empty(C, dtype=x.dtype, device=x.device)
        sc = scratch(L.pdl_bn_scratch_floats(M, C) * 4, x.device, "bn")
        _ok(L.pdl_bn_bwd(_p(x), _p(dy), _p(mask), _p(weight), _p(saved), _p(dx), _p(dres), _p(dgamma), _p(dbeta), _p(sc), M, C, int(ctx.relu),
                         _stream(x)), "batch norm backward")
        return dx, dres, dgamma, dbeta, None, None, None, None, None, None, None


def batch_norm(x, weight, bias, running_mean, running_var, training, momentum, eps, relu=False, residual=None, num_batches_tracked=None,
               partial=None):
    """``num_batches_tracked`` (int64 tensor or None) is incremented by the training forward, on the device."""
    C = x.shape[1]
    if C % 4 != 0:
        if training and num_batches_tracked is not None:
            num_batches_tracked.add_(1)
        return _torch_form(x, weight, bias, running_mean, running_var, training, momentum, eps, relu, residual)
    if training:
        if partial is not None and (partial.dim() != 3 or partial.shape[1] != C or not x.is_contiguous(memory_format=_CL)):
            partial = None  # not the tensor the statistics were computed on
        return _BatchNormAct.apply(x, residual, weight, bias, running_mean, running_var, num_batches_tracked, momentum, eps, relu, partial)
    # eval mode: an affine map of the input; gradients (if any) flow through plain tensor ops
    if torch.is_grad_enabled() and (x.requires_grad or weight.requires_grad or (residual is not None and residual.requires_grad)):
        return _torch_form(x, weight, bias, running_mean, running_var, False, momentum, eps, relu, residual)
    L = _lib.cuda()
    x = _nhwc(x)
    if residual is not None:
        residual = _nhwc(residual)
    N, _, H, W = x.shape
    y = torch.empty_like(x, memory_format=_CL)
    sc = scratch(L.pdl_bn_scratch_floats(N * H * W, C) * 4, x.device, "bn")
    _ok(L.pdl_bn_fwd_eval(_p(x), _p(residual), _p(weight), _p(bias), _p(running_mean), _p(running_var), float(eps), _p(y), _p(sc), N * H * W, C,
                          int(relu), _stream(x)), "batch norm (eval)")
    return y


def _torch_form(x, weight, bias, running_mean, running_var, training, momentum, eps, relu, residual):
    y = torch.nn.functional.batch_norm(x, running_mean, running_var, weight, bias, training, momentum, eps)
    if residual is not None:
        y = y + residual
    return torch.relu(y) if relu else y
