"""Stand-alone multi-tensor SGD launch (csrc/cuda/sgd.cu): one kernel for a whole parameter list."""
from __future__ import annotations

import ctypes

import torch

from .. import _lib

_tables = {}
_captured = {}  # tables built while a CUDA graph was being captured: the graph replays their H2D copies, so they never expire


def _table_for(params, grads, bufs):
    key = tuple((p.data_ptr(), g.data_ptr(), 0 if b is None else b.data_ptr(), p.numel()) for p, g, b in zip(params, grads, bufs))
    t = _tables.get(key) or _captured.get(key)
    if t is None:
        capturing = torch.cuda.is_current_stream_capturing()
        if len(_tables) > 64:
            _tables.clear()
        chunk = _lib.cuda().pdl_sgd_chunk_elems()
        rows, prefix, total = [], [0], 0
        for ptr_p, ptr_g, ptr_m, n in key:
            rows.append([ptr_p, ptr_g, ptr_m, n])
            total += (n + chunk - 1) // chunk
            prefix.append(total)
        dev = params[0].device
        # pinned staging + async copy: legal under stream capture (the copy becomes a graph node that re-reads the
        # pinned rows on every replay, which is why they are kept alive next to the device copies)
        host_rows = torch.tensor(rows, dtype=torch.int64).pin_memory()
        host_prefix = torch.tensor(prefix, dtype=torch.int32).pin_memory()
        t = (host_rows.to(dev, non_blocking=True), host_prefix.to(dev, non_blocking=True), total, host_rows, host_prefix)
        (_captured if capturing else _tables)[key] = t
    return t[:3]


def multi_tensor_sgd(params, grads, bufs, first_flags, lr, momentum, dampening, weight_decay, nesterov):
    """p -= lr * f(g) for every tensor in one launch per (first-step / later-step) class."""
    for want_first in (True, False):
        sel = [i for i, f in enumerate(first_flags) if bool(f) == want_first]
        if not sel:
            continue
        ps = [params[i] for i in sel]
        gs = [grads[i] if grads[i].is_contiguous() else grads[i].contiguous() for i in sel]
        bs = [bufs[i] for i in sel]
        if any(not p.is_contiguous() for p in ps):
            raise RuntimeError("native SGD needs contiguous parameters")
        table, prefix, total = _table_for(ps, gs, bs)
        stream = ctypes.c_void_p(torch.cuda.current_stream(ps[0].device).cuda_stream)
        rc = _lib.cuda().pdl_multi_tensor_sgd(ctypes.c_void_p(table.data_ptr()), ctypes.c_void_p(prefix.data_ptr()), len(ps), total, float(lr),
                                              float(momentum), float(dampening), float(weight_decay), int(bool(nesterov)), int(want_first), stream)
        if rc != 0:
            raise RuntimeError("multi-tensor SGD failed: %s" % _lib.cuda().pdl_last_error().decode())
