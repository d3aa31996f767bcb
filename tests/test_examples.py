"""The four example scripts (counterparts of the reference's four scripts) run end to end on CPU, world size 2."""
import os
import subprocess
import sys

import pytest

from conftest import free_port

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, CUDA_VISIBLE_DEVICES="", OMP_NUM_THREADS="2")


def _run(cmd, timeout=420):
    p = subprocess.run(cmd, cwd=ROOT, env=ENV, timeout=timeout, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 0, p.stdout[-3000:]
    return p.stdout


def test_convnet_spawn_example():
    out = _run([sys.executable, "examples/convnet_spawn.py", "-n", "1", "-g", "2", "-nr", "0", "--epochs", "1", "--max-steps", "3",
                "--master-port", str(free_port())])
    assert "My rank is 0" in out and "My rank is 1" in out and "Training complete in:" in out


def test_convnet_launch_example(tmp_path):
    ck = tmp_path / "ck.pt"
    out = _run([sys.executable, "-m", "pytorch_distributed_learning_b200.launch", "--nproc_per_node=2", "--master_port=%d" % free_port(),
                "examples/convnet_launch.py", "--epochs", "1", "--max-steps", "3", "--save", str(ck)])
    assert "Total step:  300" in out and "Training complete in:" in out
    import torch

    assert list(torch.load(ck).keys())[0] == "conv1.weight"


def test_restart_resumes_from_the_last_checkpoint(tmp_path):
    """launch --max_restarts + periodic training-state checkpoints: rank 1 is killed after step 5 of the first attempt, the
    group restarts, resumes at step 4 and ends with the weights of an uninterrupted run."""
    import torch

    def launch(*flags, restarts=0):
        return _run([sys.executable, "-m", "pytorch_distributed_learning_b200.launch", "--nproc_per_node=2", "--max_restarts=%d" % restarts,
                     "--master_port=%d" % free_port(), "examples/convnet_launch.py", "--epochs", "1", "--max-steps", "8"] + list(flags))

    launch("--save", str(tmp_path / "straight.pt"))
    out = launch("--save", str(tmp_path / "resumed.pt"), "--checkpoint", str(tmp_path / "train.ck"), "--checkpoint-every", "2",
                 "--inject-failure-at", "5", restarts=1)
    assert "restarting the worker group (attempt 1 of 1)" in out and "Resumed from" in out and "at step 4 (restart 1)" in out
    a, b = torch.load(tmp_path / "straight.pt"), torch.load(tmp_path / "resumed.pt")
    for k in a:
        torch.testing.assert_close(a[k], b[k], rtol=1e-5, atol=1e-7, msg=lambda m, k=k: "%s: %s" % (k, m))
    ck = torch.load(tmp_path / "train.ck", weights_only=False)
    assert ck["extra"]["step"] == 8 and "optimizer" in ck and list(ck["model"].keys())[0] == "conv1.weight"


@pytest.mark.parametrize("script", ["resnet_spawn"])
def test_resnet_spawn_example(script):
    out = _run([sys.executable, "examples/%s.py" % script, "--ngpus_per_node", "2", "--dist-url", "tcp://127.0.0.1:%d" % free_port(),
                "--max-steps", "1"], timeout=600)
    assert "[init] == local rank: 1, global rank: 1 ==" in out and "Training Finished" in out
