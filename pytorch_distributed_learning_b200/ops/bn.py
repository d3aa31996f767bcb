"""BatchNorm2d on the native kernels (csrc/cuda/bn.cu) with the ResNet block's ReLU and residual add
fused in: ``y = act(bn(x) [+ residual])``.  Training-mode batch statistics + running-stat update,
eval-mode affine, and the closed-form backward (which also returns the skip connection's gradient).
NHWC / channels-last fp32.

Used by ``nn.BatchNorm2d`` for the ResNets the reference trains (example_launch.py:26)."""
from __future__ import annotations

import torch

from .. import _lib
from .functional import _nhwc, _ok, _p, _stream, grad_buffer, scratch

_CL = torch.channels_last


class _BatchNormAct(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, residual, weight, bias, running_mean, running_var, num_batches_tracked, momentum, eps, relu, partial=None, dual=False):
        """``dual``: return the output twice (two tensor objects over the same memory).  A residual block hands one to the next
        block's first convolution and the other to its skip connection; their gradients then arrive here as two tensors and
        the backward kernel adds them while it reads them, instead of autograd summing them in a pass of its own."""
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
        ctx.bias_ref = bias.detach()
        if dual:
            ctx.set_materialize_grads(False)
            return y, y.view_as(y)
        return y

    @staticmethod
    def backward(ctx, dy, dy2=None):
        L = _lib.cuda()
        x, weight, saved, mask = ctx.saved_tensors
        if dy is None:
            dy, dy2 = dy2, None
        if dy is None:
            return (None,) * 12
        dy = _nhwc(dy)
        if dy2 is not None:
            dy2 = _nhwc(dy2)
        N, C, H, W = x.shape
        M = N * H * W
        dx = torch.empty_like(x, memory_format=_CL)
        dres = torch.empty_like(x, memory_format=_CL) if ctx.has_res else None
        dgamma = grad_buffer(weight)   # straight into DDP's gradient arena when one is installed
        dbeta = grad_buffer(ctx.bias_ref)
        sc = scratch(L.pdl_bn_scratch_floats(M, C) * 4, x.device, "bn")
        _ok(L.pdl_bn_bwd2(_p(x), _p(dy), _p(dy2), _p(mask), _p(weight), _p(saved), _p(dx), _p(dres), _p(dgamma), _p(dbeta), _p(sc), M, C,
                          int(ctx.relu), _stream(x)), "batch norm backward")
        return dx, dres, dgamma, dbeta, None, None, None, None, None, None, None, None


def batch_norm(x, weight, bias, running_mean, running_var, training, momentum, eps, relu=False, residual=None, num_batches_tracked=None,
               partial=None, dual=False):
    """``num_batches_tracked`` (int64 tensor or None) is incremented by the training forward, on the device.  ``dual=True``
    returns the output as a pair of aliases when the fused training kernel runs (see ``_BatchNormAct.forward``), else a
    pair holding the same tensor twice."""
    if dual:
        C = x.shape[1]
        if training and C % 4 == 0:
            if partial is not None and (partial.dim() != 3 or partial.shape[1] != C or not x.is_contiguous(memory_format=_CL)):
                partial = None
            return _BatchNormAct.apply(x, residual, weight, bias, running_mean, running_var, num_batches_tracked, momentum, eps, relu, partial, True)
        y = batch_norm(x, weight, bias, running_mean, running_var, training, momentum, eps, relu, residual, num_batches_tracked, partial)
        return y, y
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
