"""In-tree native build: g++ for the host runtime, nvcc (sm_100a) for the kernels.

The libraries are plain C-ABI shared objects loaded with ctypes (no torch headers, so a full
rebuild takes seconds and cross-compiles on a box without a GPU).  Outputs live in
``pytorch_distributed_learning_b200/_native/`` so they travel with the source tree.
"""
from __future__ import annotations

import glob
import os
import shutil
import subprocess
import sys

PKG_DIR = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(PKG_DIR)
NATIVE_DIR = os.path.join(PKG_DIR, "_native")
HOST_LIB = os.path.join(NATIVE_DIR, "libpdl_host.so")
CUDA_LIB = os.path.join(NATIVE_DIR, "libpdl_cuda.so")

HOST_SRC = sorted(glob.glob(os.path.join(ROOT, "csrc", "host", "*.cpp")))
HOST_HDR = sorted(glob.glob(os.path.join(ROOT, "csrc", "host", "*.h")))


def _cuda_sources():
    return sorted(glob.glob(os.path.join(ROOT, "csrc", "cuda", "*.cu")))


def _cuda_headers():
    return sorted(glob.glob(os.path.join(ROOT, "csrc", "cuda", "*.cuh")) + glob.glob(os.path.join(ROOT, "include", "*.h")))


def _stale(target, deps):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(d) > t for d in deps if os.path.exists(d))


def _run(cmd):
    proc = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    if proc.returncode != 0:
        raise RuntimeError("native build failed:\n$ %s\n%s" % (" ".join(cmd), proc.stdout))
    return proc.stdout


def build_host(force=False, verbose=False):
    """Compile csrc/host/*.cpp -> libpdl_host.so (store, CPU collectives, bucket assignment)."""
    os.makedirs(NATIVE_DIR, exist_ok=True)
    if not force and not _stale(HOST_LIB, HOST_SRC + HOST_HDR):
        return HOST_LIB
    if not HOST_SRC:
        raise RuntimeError("csrc/host sources are missing and %s is not built" % HOST_LIB)
    tmp = HOST_LIB + ".tmp.%d" % os.getpid()
    out = _run(["g++", "-O2", "-std=c++17", "-fPIC", "-shared", "-pthread", "-o", tmp] + HOST_SRC)
    os.replace(tmp, HOST_LIB)
    if verbose:
        print(out)
    return HOST_LIB


def nvcc_path():
    for cand in (os.environ.get("NVCC"), shutil.which("nvcc"), "/usr/local/cuda/bin/nvcc"):
        if cand and os.path.exists(cand):
            return cand
    return None


def build_cuda(force=False, verbose=False):
    """Compile csrc/cuda/*.cu -> libpdl_cuda.so for sm_100a only (no other arch, no PTX fallback)."""
    os.makedirs(NATIVE_DIR, exist_ok=True)
    srcs = _cuda_sources()
    if not force and not _stale(CUDA_LIB, srcs + _cuda_headers()):
        return CUDA_LIB
    nvcc = nvcc_path()
    if nvcc is None or not srcs:
        raise RuntimeError("cannot build %s: nvcc=%r, %d sources" % (CUDA_LIB, nvcc, len(srcs)))
    objs = []
    obj_dir = os.path.join(ROOT, "build", "cuda")
    os.makedirs(obj_dir, exist_ok=True)
    procs = []
    for s in srcs:
        o = os.path.join(obj_dir, os.path.basename(s) + ".o")
        objs.append(o)
        if force or _stale(o, [s] + _cuda_headers()):
            cmd = [nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-O3", "-std=c++17", "-lineinfo",
                   "-Xcompiler", "-fPIC", "-Xptxas", "-v" if verbose else "-warn-spills",
                   "-I", os.path.join(ROOT, "include"), "-I", os.path.join(ROOT, "csrc", "cuda"), "-c", s, "-o", o]
            procs.append((cmd, subprocess.Popen(cmd, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)))
    for cmd, p in procs:
        out, _ = p.communicate()
        if p.returncode != 0:
            raise RuntimeError("native build failed:\n$ %s\n%s" % (" ".join(cmd), out))
        if verbose:
            print(out)
    tmp = CUDA_LIB + ".tmp.%d" % os.getpid()
    _run([nvcc, "-gencode", "arch=compute_100a,code=sm_100a", "-shared", "-Xcompiler", "-fPIC", "-o", tmp] + objs +
         ["-lcuda", "-lcudart_static", "-lpthread", "-ldl", "-lrt"])
    os.replace(tmp, CUDA_LIB)
    return CUDA_LIB


def build_all(force=False, verbose=False):
    return build_host(force, verbose), build_cuda(force, verbose)


if __name__ == "__main__":
    force = "--force" in sys.argv
    verbose = "-v" in sys.argv
    print(build_host(force, verbose))
    print(build_cuda(force, verbose))
