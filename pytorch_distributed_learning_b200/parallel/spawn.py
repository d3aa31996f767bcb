"""``spawn(fn, args, nprocs)``: one child process per local GPU, ``fn(i, *args)`` in each.

Same call shape the reference uses (mpspawn_dist.py:140 ``mp.spawn(train, nprocs=args.gpus,
args=(args,))``, example_mp.py:27).  Failure handling: a child that raises writes its
traceback to a private temp file; the parent notices the non-zero exit, terminates the
siblings (SIGTERM, then SIGKILL after a grace period) and re-raises with the child's
traceback.  Children die with the parent (PR_SET_PDEATHSIG on Linux).
"""
from __future__ import annotations

import multiprocessing
import multiprocessing.connection
import os
import pickle
import signal
import sys
import tempfile
import time
import traceback


class ProcessException(Exception):
    def __init__(self, msg, error_index, pid):
        super().__init__(msg)
        self.msg, self.error_index, self.pid = msg, error_index, pid

    def __reduce__(self):
        return type(self), (self.msg, self.error_index, self.pid)


class ProcessRaisedException(ProcessException):
    """A child raised a Python exception (its traceback is in the message)."""


class ProcessExitedException(ProcessException):
    """A child exited with a non-zero code or was killed by a signal."""

    def __init__(self, msg, error_index, pid, exit_code, signal_name=None):
        super().__init__(msg, error_index, pid)
        self.exit_code, self.signal_name = exit_code, signal_name

    def __reduce__(self):
        return type(self), (self.msg, self.error_index, self.pid, self.exit_code, self.signal_name)


def _die_with_parent():
    if sys.platform.startswith("linux"):
        try:
            import ctypes

            ctypes.CDLL("libc.so.6", use_errno=True).prctl(1, signal.SIGTERM)  # PR_SET_PDEATHSIG
        except Exception:
            pass


def _child_main(fn, index, args, error_path):
    _die_with_parent()
    try:
        fn(index, *args)
    except KeyboardInterrupt:
        pass
    except Exception:
        with open(error_path, "wb") as fh:
            pickle.dump(traceback.format_exc(), fh)
        sys.exit(1)


class ProcessContext:
    def __init__(self, processes, error_paths):
        self.processes, self.error_paths = processes, error_paths
        self.sentinels = {p.sentinel: i for i, p in enumerate(processes)}

    def pids(self):
        return [int(p.pid) for p in self.processes]

    def _reap_siblings(self, grace_s=30.0):
        for p in self.processes:
            if p.is_alive():
                p.terminate()
        deadline = time.monotonic() + grace_s
        for p in self.processes:
            p.join(max(0.0, deadline - time.monotonic()))
        for p in self.processes:
            if p.is_alive():
                p.kill()
                p.join()

    def join(self, timeout=None, grace_period=30.0):
        """True once every child has exited cleanly; raises on the first failure."""
        if not self.sentinels:
            return True
        ready = multiprocessing.connection.wait(self.sentinels.keys(), timeout=timeout)
        failed = None
        for s in ready:
            i = self.sentinels.pop(s)
            p = self.processes[i]
            p.join()
            if p.exitcode != 0 and failed is None:
                failed = i
        if failed is None:
            return not self.sentinels
        self._reap_siblings(grace_period)
        p = self.processes[failed]
        path = self.error_paths[failed]
        if os.path.exists(path) and os.path.getsize(path) > 0:
            with open(path, "rb") as fh:
                tb = pickle.load(fh)
            raise ProcessRaisedException("\n\n-- Process %d terminated with the following error:\n%s" % (failed, tb), failed, p.pid)
        if p.exitcode < 0:
            try:
                name = signal.Signals(-p.exitcode).name
            except ValueError:
                name = "<unknown %d>" % -p.exitcode
            raise ProcessExitedException("process %d terminated with signal %s" % (failed, name), failed, p.pid, p.exitcode, name)
        raise ProcessExitedException("process %d terminated with exit code %d" % (failed, p.exitcode), failed, p.pid, p.exitcode)


def start_processes(fn, args=(), nprocs=1, join=True, daemon=False, start_method="spawn"):
    ctx = multiprocessing.get_context(start_method)
    procs, paths = [], []
    tmpdir = tempfile.mkdtemp(prefix="pdl_spawn_")
    for i in range(nprocs):
        path = os.path.join(tmpdir, "%d.err" % i)
        p = ctx.Process(target=_child_main, args=(fn, i, args, path), daemon=daemon)
        p.start()
        procs.append(p)
        paths.append(path)
    context = ProcessContext(procs, paths)
    if not join:
        return context
    try:
        while not context.join():
            pass
    finally:
        for path in paths:
            if os.path.exists(path):
                os.unlink(path)
        try:
            os.rmdir(tmpdir)
        except OSError:
            pass
    return None


def spawn(fn, args=(), nprocs=1, join=True, daemon=False, start_method="spawn"):
    if start_method != "spawn":
        raise ValueError("spawn() only supports start_method='spawn' (CUDA cannot be re-initialised after fork)")
    return start_processes(fn, args, nprocs, join, daemon, start_method="spawn")
