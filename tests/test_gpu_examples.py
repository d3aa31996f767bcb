"""The launcher examples on a real GPU (world size 1): eager and --graphed loops print the same kind of lines."""
import os
import re
import subprocess
import sys

import pytest

from conftest import free_port

pytestmark = pytest.mark.gpu

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _launch(script, *flags):
    cmd = [sys.executable, "-m", "pytorch_distributed_learning_b200.launch", "--nproc_per_node=1", "--master_port=%d" % free_port(),
           "examples/%s.py" % script] + list(flags)
    p = subprocess.run(cmd, cwd=ROOT, env=dict(os.environ, CUDA_VISIBLE_DEVICES="0"), timeout=300, stdout=subprocess.PIPE,
                       stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 0, p.stdout[-3000:]
    return p.stdout


@pytest.mark.parametrize("graphed", [False, True])
def test_resnet_launch_example_on_gpu(graphed):
    out = _launch("resnet_launch", "--max-steps", "26", *(["--graphed"] if graphed else []))
    m = re.search(r"step: \[ 25/\d+\] \[1/5\] \| loss: ([0-9.]+) \| acc:\s+([0-9.]+)%", out)
    assert m, out[-2000:]
    assert 0.5 < float(m.group(1)) < 10.0 and 0.0 <= float(m.group(2)) <= 100.0
    assert "Training Finished" in out


def _gpus():
    import torch

    return torch.cuda.device_count()


@pytest.mark.skipif("_gpus() < 2", reason="needs 2 GPUs")
def test_examples_on_two_gpus():
    env = dict(os.environ, CUDA_VISIBLE_DEVICES="0,1")
    cmd = [sys.executable, "-m", "pytorch_distributed_learning_b200.launch", "--nproc_per_node=2", "--master_port=%d" % free_port(),
           "examples/convnet_launch.py", "--epochs", "1", "--max-steps", "120", "--graphed"]
    p = subprocess.run(cmd, cwd=ROOT, env=env, timeout=300, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 0, p.stdout[-3000:]
    assert re.search(r"Epoch \[1/1\], Step \[100/300\], Loss: [0-9.]+", p.stdout), p.stdout[-2000:]  # 60000 / (2 ranks x 100)
    cmd = [sys.executable, "examples/resnet_spawn.py", "--ngpus_per_node", "2", "--dist-url", "tcp://127.0.0.1:%d" % free_port(),
           "--max-steps", "26", "--graphed"]
    p = subprocess.run(cmd, cwd=ROOT, env=env, timeout=300, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True)
    assert p.returncode == 0, p.stdout[-3000:]
    assert re.search(r"step: \[ 25/196\] \[1/5\] \| loss: [0-9.]+ \| acc:", p.stdout), p.stdout[-2000:]  # 50000 / (2 x 128)
    assert "[init] == local rank: 1, global rank: 1 ==" in p.stdout


def test_convnet_launch_example_graphed_on_gpu():
    out = _launch("convnet_launch", "--epochs", "1", "--max-steps", "120", "--graphed")
    assert re.search(r"Epoch \[1/1\], Step \[100/\d+\], Loss: [0-9.]+", out), out[-2000:]
