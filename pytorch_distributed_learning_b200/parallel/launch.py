"""``python -m pytorch_distributed_learning_b200.launch --nproc_per_node=G ... script.py [args]``

Stand-in for the launcher the reference documents (README.md:341-343):
``--nproc_per_node --nnodes --node_rank --master_addr --master_port [--use_env]``.
Each worker gets RANK, LOCAL_RANK, WORLD_SIZE, LOCAL_WORLD_SIZE, GROUP_RANK, MASTER_ADDR,
MASTER_PORT and OMP_NUM_THREADS in its environment (what launch_dist.py:45-46 reads) and,
unless ``--use_env`` is given, ``--local-rank=N`` appended to its argv.  Workers are polled
every 100 ms; the first non-zero exit terminates the rest and becomes the launcher's exit
code, unless ``--max_restarts=N`` asks for the whole group to be started again (default 0, as in torch).
"""
from __future__ import annotations

import argparse
import os
import signal
import subprocess
import sys
import time


def build_parser():
    p = argparse.ArgumentParser(prog="pytorch_distributed_learning_b200.launch", description=__doc__,
                                formatter_class=argparse.RawDescriptionHelpFormatter)
    p.add_argument("--nnodes", type=int, default=1)
    p.add_argument("--node_rank", "--node-rank", type=int, default=0)
    p.add_argument("--nproc_per_node", "--nproc-per-node", type=str, default="1", help="int, or 'gpu' for one per visible GPU")
    p.add_argument("--master_addr", "--master-addr", type=str, default="127.0.0.1")
    p.add_argument("--master_port", "--master-port", type=int, default=29500)
    p.add_argument("--use_env", "--use-env", action="store_true", help="do not append --local-rank=N to the script's argv")
    p.add_argument("--max_restarts", "--max-restarts", type=int, default=0, help="restart the whole worker group this many times after a failure (single node)")
    p.add_argument("-m", "--module", action="store_true", help="run the script as a python module")
    p.add_argument("--no_python", "--no-python", action="store_true", help="exec the script directly")
    p.add_argument("training_script", type=str)
    p.add_argument("training_script_args", nargs=argparse.REMAINDER)
    return p


def _nproc(value):
    if value in ("gpu", "auto"):
        import torch

        n = torch.cuda.device_count()
        if n < 1:
            raise SystemExit("--nproc_per_node=%s but no GPU is visible" % value)
        return n
    return int(value)


def worker_env(base, node_rank, local_rank, nproc, nnodes, master_addr, master_port):
    env = dict(base)
    world = nproc * nnodes
    env.update({
        "RANK": str(node_rank * nproc + local_rank),
        "LOCAL_RANK": str(local_rank),
        "WORLD_SIZE": str(world),
        "LOCAL_WORLD_SIZE": str(nproc),
        "GROUP_RANK": str(node_rank),
        "MASTER_ADDR": master_addr,
        "MASTER_PORT": str(master_port),
    })
    if "OMP_NUM_THREADS" not in env and nproc > 1:
        env["OMP_NUM_THREADS"] = "1"
    return env


def worker_cmd(args, local_rank):
    cmd = []
    if not args.no_python:
        cmd += [sys.executable, "-u"]
        if args.module:
            cmd.append("-m")
    cmd.append(args.training_script)
    cmd += list(args.training_script_args)
    if not args.use_env:
        cmd.append("--local-rank=%d" % local_rank)
    return cmd


def run(args):
    """Run the worker group; after a failure, start the WHOLE group again up to ``--max_restarts`` times (what the elastic
    agent behind ``torch.distributed.launch`` does; default 0 = fail fast, [T] torch/distributed/run.py:491-505).  Restarted
    workers see TORCHELASTIC_RESTART_COUNT = 1, 2, ... and are expected to resume from their last checkpoint."""
    nproc = _nproc(args.nproc_per_node)
    max_restarts = max(0, int(getattr(args, "max_restarts", 0) or 0))
    if max_restarts and args.nnodes > 1:
        raise SystemExit("--max_restarts needs a re-rendezvous across nodes, which env:// / tcp:// cannot do; use it with --nnodes=1")
    attempt = 0
    while True:
        code = _run_group(args, nproc, attempt, max_restarts)
        if code == 0 or attempt >= max_restarts:  # (a signal to the launcher itself exits through on_signal, not here)
            return code
        attempt += 1
        sys.stderr.write("launch: restarting the worker group (attempt %d of %d)\n" % (attempt, max_restarts))


def _run_group(args, nproc, restart_count, max_restarts):
    procs = []
    for lr in range(nproc):
        env = worker_env(os.environ, args.node_rank, lr, nproc, args.nnodes, args.master_addr, args.master_port)
        env["TORCHELASTIC_RESTART_COUNT"], env["TORCHELASTIC_MAX_RESTARTS"] = str(restart_count), str(max_restarts)
        procs.append(subprocess.Popen(worker_cmd(args, lr), env=env))

    def terminate_all(sig=signal.SIGTERM):
        for p in procs:
            if p.poll() is None:
                try:
                    p.send_signal(sig)
                except OSError:
                    pass

    def on_signal(signum, _frame):
        terminate_all(signum)
        raise SystemExit(128 + signum)

    signal.signal(signal.SIGINT, on_signal)
    signal.signal(signal.SIGTERM, on_signal)

    exit_code = 0
    alive = set(range(nproc))
    while alive:
        for i in list(alive):
            rc = procs[i].poll()
            if rc is None:
                continue
            alive.discard(i)
            if rc != 0 and exit_code == 0:
                exit_code = rc if rc > 0 else 128 - rc
                sys.stderr.write("launch: worker %d (pid %d) failed with exit code %d; stopping the others\n" % (i, procs[i].pid, rc))
                terminate_all()
                deadline = time.time() + 30
                for p in procs:
                    try:
                        p.wait(max(0.0, deadline - time.time()))
                    except subprocess.TimeoutExpired:
                        p.kill()
                alive.clear()
        time.sleep(0.1)
    return exit_code


def main(argv=None):
    args = build_parser().parse_args(argv)
    sys.exit(run(args))


if __name__ == "__main__":
    main()
