"""Timeline annotations: NVTX ranges around the framework's host-side phases, plus an optional host trace.

torch's DDP marks its forward with ``record_function("DistributedDataParallel.forward")``
([T] torch/nn/parallel/distributed.py:1694); the reference itself has no tracing.  Here:

* ``PDL_NVTX=1`` -- ``trace_range(name)`` pushes/pops an NVTX range (visible in Nsight Systems / CUPTI-based tools)
  around ``DistributedDataParallel.forward``, every bucket launch (with its index, bytes and algorithm) and every
  ``TrainStep`` call.  Off by default: the calls cost ~1 us each and the ConvNet step is 120 us.
* ``PDL_TRACE=/path/prefix`` -- the same ranges are also recorded on the host (``perf_counter_ns``) and written at exit as
  Chrome-trace JSON, one file per rank (``<prefix>.rank<r>.json``; open in chrome://tracing or Perfetto).  Works without
  a GPU, which is how the CPU tests exercise it.
"""
from __future__ import annotations

import atexit
import contextlib
import json
import os
import threading
import time

_lock = threading.Lock()
_events = []
_state = {"nvtx": None, "path": None, "registered": False}


def _nvtx_enabled():
    if _state["nvtx"] is None:
        on = os.environ.get("PDL_NVTX", "0") == "1"
        if on:
            try:
                import torch

                on = torch.cuda.is_available()
            except Exception:  # pragma: no cover
                on = False
        _state["nvtx"] = on
    return _state["nvtx"]


def _trace_path():
    if _state["path"] is None:
        _state["path"] = os.environ.get("PDL_TRACE", "")
        if _state["path"] and not _state["registered"]:
            _state["registered"] = True
            atexit.register(flush)
    return _state["path"]


def reset():
    """Re-read the environment switches (tests flip them at run time)."""
    _state["nvtx"] = _state["path"] = None


def enabled():
    return _nvtx_enabled() or bool(_trace_path())


@contextlib.contextmanager
def trace_range(name, **args):
    """``with trace_range("pdl.bucket", index=0, bytes=454696): ...`` -- no-op unless PDL_NVTX / PDL_TRACE is set."""
    nvtx, path = _nvtx_enabled(), _trace_path()
    if not nvtx and not path:
        yield
        return
    if nvtx:
        import torch

        torch.cuda.nvtx.range_push(name if not args else "%s %s" % (name, " ".join("%s=%s" % kv for kv in args.items())))
    t0 = time.perf_counter_ns() if path else 0
    try:
        yield
    finally:
        if path:
            t1 = time.perf_counter_ns()
            with _lock:
                _events.append({"name": name, "ph": "X", "ts": t0 / 1e3, "dur": (t1 - t0) / 1e3, "pid": os.getpid(),
                                "tid": threading.get_ident() & 0xFFFF, "args": args})
        if nvtx:
            torch.cuda.nvtx.range_pop()


def events():
    with _lock:
        return list(_events)


def flush(path=None):
    """Write the recorded host ranges as Chrome-trace JSON; returns the file name (or None if nothing was recorded)."""
    prefix = path or _trace_path()
    with _lock:
        evs = list(_events)
    if not prefix or not evs:
        return None
    rank = os.environ.get("RANK", "0")
    name = "%s.rank%s.json" % (prefix, rank)
    with open(name, "w") as f:
        json.dump({"traceEvents": evs, "displayTimeUnit": "ms"}, f)
    return name
