"""Python face of the C++ reducer core (csrc/host/reducer.cpp).

torch keeps this bookkeeping in its C++ ``Reducer`` ([T] torch/include/torch/csrc/distributed/c10d/reducer.hpp:74,231,286,
380-382; built by the DDP wrap at launch_dist.py:61): per-iteration ready flags, per-bucket pending counts, strictly
in-order bucket launches, the tail flush at the end of backward and the gradient-arrival order of the first iteration.
``DistributedDataParallel`` (ddp.py) asks this object WHICH buckets to launch; it owns the tensors and the kernels.
"""
from __future__ import annotations

import ctypes

from .. import _lib


class ReducerCore:
    def __init__(self, num_params, bucket_of):
        """``bucket_of[i]`` = bucket index of parameter i."""
        self._L = _lib.host()
        self._h = None
        self._out = None
        self.set_layout(num_params, bucket_of)

    def set_layout(self, num_params, bucket_of):
        nb = max(bucket_of) + 1 if num_params else 0
        arr = (ctypes.c_int32 * max(num_params, 1))(*[int(b) for b in bucket_of])
        if self._h is None:
            self._h = self._L.pdl_reducer_create(num_params, arr, nb)
            if not self._h:
                raise ValueError("reducer: invalid bucket layout (%d parameters, %d buckets)" % (num_params, nb))
        elif self._L.pdl_reducer_set_layout(self._h, num_params, arr, nb) != 0:
            raise ValueError("reducer: invalid bucket layout (%d parameters, %d buckets)" % (num_params, nb))
        self.num_params, self.num_buckets = num_params, nb
        self._out = (ctypes.c_int32 * max(nb, num_params, 1))()

    def prepare(self, record_arrival=False):
        self._L.pdl_reducer_prepare(self._h, 1 if record_arrival else 0)

    def mark_ready(self, param):
        """Buckets that may be launched now, in launch order.  Raises on a second gradient for ``param`` in one iteration."""
        n = self._L.pdl_reducer_mark_ready(self._h, int(param), self._out, self.num_buckets)
        if n == 0:
            return ()
        if n == -2:
            raise RuntimeError("duplicate")
        if n < 0:
            raise IndexError("reducer: parameter index %d (code %d)" % (param, n))
        return self._out[:n]

    def finalize(self):
        """Buckets still to be launched at the end of backward, in order."""
        n = self._L.pdl_reducer_finalize(self._h, self._out, self.num_buckets)
        if n < 0:
            raise RuntimeError("reducer: finalize failed (%d)" % n)
        return self._out[:n]

    def is_ready(self, param):
        return bool(self._L.pdl_reducer_is_ready(self._h, int(param)))

    @property
    def next_bucket(self):
        return self._L.pdl_reducer_next_bucket(self._h)

    def arrival(self):
        n = self._L.pdl_reducer_arrival(self._h, self._out, len(self._out))
        return list(self._out[:n])

    def close(self):
        if self._h:
            self._L.pdl_reducer_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:  # interpreter shutdown
            pass
