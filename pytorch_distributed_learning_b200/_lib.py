"""ctypes bindings for the two in-tree native libraries.

``host()`` -> libpdl_host.so  (C++: TCP store, CPU collectives, bucket assignment; no CUDA)
``cuda()`` -> libpdl_cuda.so  (sm_100a kernels + symmetric-memory runtime; needs a GPU)

The host library is rebuilt on demand with g++ (seconds).  The CUDA library must already be
built (``python -m pytorch_distributed_learning_b200._build`` or ``__graft_entry__.build()``);
on a GPU box a missing library is a hard error, never a silent PyTorch fallback.
"""
from __future__ import annotations

import ctypes
import os
import threading
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_uint64, c_void_p

from . import _build

_lock = threading.Lock()
_host = None
_cuda = None


def host():
    global _host
    if _host is None:
        with _lock:
            if _host is None:
                # PDL_HOST_LIB: load an alternative build of the host runtime (scripts/sanitize_host.sh uses it for the
                # AddressSanitizer / ThreadSanitizer variants)
                path = os.environ.get("PDL_HOST_LIB") or _build.build_host()
                lib = ctypes.CDLL(path)
                _declare_host(lib)
                _host = lib
    return _host


def _declare_host(L):
    L.pdl_store_server_start.restype = c_void_p
    L.pdl_store_server_start.argtypes = [c_char_p, c_int, POINTER(c_int)]
    L.pdl_store_server_stop.argtypes = [c_void_p]
    L.pdl_store_client_connect.restype = c_void_p
    L.pdl_store_client_connect.argtypes = [c_char_p, c_int, c_double]
    L.pdl_store_client_close.argtypes = [c_void_p]
    L.pdl_store_set.argtypes = [c_void_p, c_char_p, c_void_p, c_uint64]
    L.pdl_store_get.argtypes = [c_void_p, c_char_p, c_double, POINTER(c_void_p), POINTER(c_uint64)]
    L.pdl_store_add.argtypes = [c_void_p, c_char_p, c_int64, POINTER(c_int64)]
    L.pdl_store_wait.argtypes = [c_void_p, POINTER(c_char_p), c_int, c_double]
    L.pdl_store_check.argtypes = [c_void_p, POINTER(c_char_p), c_int]
    L.pdl_store_delete.argtypes = [c_void_p, c_char_p]
    L.pdl_store_num_keys.restype = c_int64
    L.pdl_store_num_keys.argtypes = [c_void_p]
    L.pdl_store_compare_set.argtypes = [c_void_p, c_char_p, c_void_p, c_uint64, c_void_p, c_uint64, POINTER(c_void_p), POINTER(c_uint64)]
    L.pdl_free.argtypes = [c_void_p]
    L.pdl_cpu_comm_create.restype = c_void_p
    L.pdl_cpu_comm_create.argtypes = [c_void_p, c_char_p, c_int, c_int, c_char_p, c_double]
    L.pdl_cpu_comm_destroy.argtypes = [c_void_p]
    L.pdl_cpu_allreduce.argtypes = [c_void_p, c_void_p, c_uint64, c_int, c_int]
    L.pdl_cpu_broadcast.argtypes = [c_void_p, c_void_p, c_uint64, c_int]
    L.pdl_cpu_allgather.argtypes = [c_void_p, c_void_p, c_void_p, c_uint64]
    L.pdl_cpu_barrier.argtypes = [c_void_p]
    L.pdl_cpu_send.argtypes = [c_void_p, c_void_p, c_uint64, c_int]
    L.pdl_cpu_recv.argtypes = [c_void_p, c_void_p, c_uint64, c_int]
    L.pdl_bucket_assign.argtypes = [POINTER(c_int64), POINTER(c_int32), c_int, POINTER(c_int64), c_int, POINTER(c_int32)]
    L.pdl_reducer_create.restype = c_void_p
    L.pdl_reducer_create.argtypes = [c_int, POINTER(c_int32), c_int]
    L.pdl_reducer_destroy.restype = None
    L.pdl_reducer_destroy.argtypes = [c_void_p]
    L.pdl_reducer_set_layout.argtypes = [c_void_p, c_int, POINTER(c_int32), c_int]
    L.pdl_reducer_prepare.restype = None
    L.pdl_reducer_prepare.argtypes = [c_void_p, c_int]
    L.pdl_reducer_mark_ready.argtypes = [c_void_p, c_int, POINTER(c_int32), c_int]
    L.pdl_reducer_finalize.argtypes = [c_void_p, POINTER(c_int32), c_int]
    L.pdl_reducer_is_ready.argtypes = [c_void_p, c_int]
    L.pdl_reducer_next_bucket.argtypes = [c_void_p]
    L.pdl_reducer_arrival.argtypes = [c_void_p, POINTER(c_int32), c_int]


def cuda_available_lib():
    return os.path.exists(_build.CUDA_LIB)


def cuda():
    """Load the sm_100a kernel library.  Raises if it has not been built."""
    global _cuda
    if _cuda is None:
        with _lock:
            if _cuda is None:
                if not os.path.exists(_build.CUDA_LIB):
                    raise RuntimeError(
                        "native CUDA library %s is missing; run `python -m pytorch_distributed_learning_b200._build` "
                        "(there is no PyTorch fallback for the GPU kernels)" % _build.CUDA_LIB
                    )
                lib = ctypes.CDLL(_build.CUDA_LIB)
                from . import _cuda_decl

                _cuda_decl.declare(lib)
                _cuda = lib
    return _cuda
