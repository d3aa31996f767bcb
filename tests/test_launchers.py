import os
import subprocess
import sys
import time

import pytest

import _workers
import pytorch_distributed_learning_b200 as pdl
from pytorch_distributed_learning_b200 import launch
from pytorch_distributed_learning_b200.parallel.spawn import ProcessExitedException, ProcessRaisedException

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_spawn_runs_every_rank(tmp_path):
    assert pdl.multiprocessing.spawn(_workers.ok_worker, args=(str(tmp_path),), nprocs=3) is None
    assert sorted(os.listdir(tmp_path)) == ["ok0", "ok1", "ok2"]


def test_spawn_propagates_child_exception_and_kills_siblings(tmp_path):
    t0 = time.time()
    with pytest.raises(ProcessRaisedException, match="boom from rank 1") as ei:
        pdl.multiprocessing.spawn(_workers.failing, args=(str(tmp_path),), nprocs=2)
    assert ei.value.error_index == 1
    assert time.time() - t0 < 45  # the sleeping sibling was terminated, not waited for
    assert not os.path.exists(tmp_path / "survivor")


def test_spawn_reports_exit_code():
    with pytest.raises(ProcessExitedException) as ei:
        pdl.multiprocessing.spawn(_workers.exiting, nprocs=2)
    assert ei.value.exit_code == 7 and ei.value.error_index == 0


def test_spawn_nonblocking_context(tmp_path):
    ctx = pdl.multiprocessing.spawn(_workers.ok_worker, args=(str(tmp_path),), nprocs=2, join=False)
    assert len(ctx.pids()) == 2
    while not ctx.join():
        pass


def test_launch_env_contract():
    env = launch.worker_env({}, node_rank=1, local_rank=2, nproc=4, nnodes=2, master_addr="10.0.0.9", master_port=8888)
    assert env == {"RANK": "6", "LOCAL_RANK": "2", "WORLD_SIZE": "8", "LOCAL_WORLD_SIZE": "4", "GROUP_RANK": "1",
                   "MASTER_ADDR": "10.0.0.9", "MASTER_PORT": "8888", "OMP_NUM_THREADS": "1"}
    args = launch.build_parser().parse_args(["--nproc_per_node=2", "--nnodes=1", "--node_rank=0", "--master_addr=127.0.0.1",
                                             "--master_port=1234", "train.py", "--epochs", "3"])
    assert launch.worker_cmd(args, 1)[-4:] == ["train.py", "--epochs", "3", "--local-rank=1"]
    args = launch.build_parser().parse_args(["--use_env", "train.py"])
    assert launch.worker_cmd(args, 0)[-1] == "train.py"


def test_launch_end_to_end(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(
        "import os,sys\n"
        "open(os.path.join(%r, 'r'+os.environ['RANK']), 'w').write(' '.join([os.environ['LOCAL_RANK'], os.environ['WORLD_SIZE'], sys.argv[-1]]))\n" % str(tmp_path))
    rc = subprocess.run([sys.executable, "-m", "pytorch_distributed_learning_b200.launch", "--nproc_per_node=2", "--master_port=29411",
                         str(script)], cwd=ROOT, timeout=120).returncode
    assert rc == 0
    assert (tmp_path / "r0").read_text() == "0 2 --local-rank=0" and (tmp_path / "r1").read_text() == "1 2 --local-rank=1"


def test_launch_failure_stops_the_job(tmp_path):
    script = tmp_path / "f.py"
    script.write_text("import os,sys,time\nif os.environ['RANK']=='1': sys.exit(3)\ntime.sleep(60)\n")
    t0 = time.time()
    rc = subprocess.run([sys.executable, "-m", "pytorch_distributed_learning_b200.launch", "--nproc_per_node=2", "--use_env", str(script)],
                        cwd=ROOT, timeout=120).returncode
    assert rc == 3 and time.time() - t0 < 45


def test_two_nodes_on_localhost(tmp_path):
    """The reference README's worked example is 2 machines x 1 process; here both "nodes" are launcher processes on
    this host (node_rank 0 and 1) that must rendezvous through node 0's store and train together."""
    import subprocess
    import sys

    from conftest import free_port

    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    port = free_port()
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2")
    procs = []
    for node in (0, 1):
        cmd = [sys.executable, "-m", "pytorch_distributed_learning_b200.launch", "--nproc_per_node=1", "--nnodes=2", "--node_rank=%d" % node,
               "--master_addr=127.0.0.1", "--master_port=%d" % port, "examples/resnet_launch.py", "--max-steps", "1"]
        procs.append(subprocess.Popen(cmd, cwd=root, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True))
    outs = []
    for p in procs:
        out, _ = p.communicate(timeout=600)
        outs.append(out)
        assert p.returncode == 0, out[-3000:]
    assert "[init] == local rank: 0, global rank: 0 ==" in outs[0] and "[init] == local rank: 0, global rank: 1 ==" in outs[1]
    assert "Training Finished" in outs[0] and "Training Finished" not in outs[1]  # only global rank 0 prints


def test_launch_max_restarts_runs_the_group_again(tmp_path):
    """--max_restarts: after a worker fails, the whole group is started again (with TORCHELASTIC_RESTART_COUNT bumped),
    rendezvouses on the same port and finishes; without attempts left the failure is final."""
    from conftest import free_port

    script = tmp_path / "flaky.py"
    script.write_text(
        "import os, sys\n"
        "sys.path.insert(0, %r)\n"
        "import torch, pytorch_distributed_learning_b200 as pdl\n"
        "attempt, rank = int(os.environ['TORCHELASTIC_RESTART_COUNT']), int(os.environ['RANK'])\n"
        "pdl.distributed.init_process_group('gloo')\n"
        "t = torch.ones(3) * (rank + 1)\n"
        "pdl.distributed.all_reduce(t)\n"
        "if attempt < int(os.environ['FAIL_UNTIL']) and rank == 1:\n"
        "    os._exit(7)\n"
        "pdl.distributed.barrier()\n"
        "print('rank %%d done on attempt %%d sum %%d' %% (rank, attempt, int(t[0])), flush=True)\n"
        "pdl.distributed.destroy_process_group()\n" % ROOT)
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="", FAIL_UNTIL="1")
    cmd = [sys.executable, "-m", "pytorch_distributed_learning_b200.launch", "--nproc_per_node=2", "--use_env", "--max_restarts=2",
           "--master_port=%d" % free_port(), str(script)]
    p = subprocess.run(cmd, cwd=ROOT, env=env, timeout=300, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 0, p.stdout[-2000:]
    assert "restarting the worker group (attempt 1 of 2)" in p.stdout
    assert "rank 0 done on attempt 1 sum 3" in p.stdout and "rank 1 done on attempt 1 sum 3" in p.stdout
    env["FAIL_UNTIL"] = "5"  # keeps failing: one restart, then the failure is final
    cmd[-2], cmd[-3] = "--master_port=%d" % free_port(), "--max_restarts=1"
    p = subprocess.run(cmd, cwd=ROOT, env=env, timeout=300, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 7 and p.stdout.count("restarting the worker group") == 1
