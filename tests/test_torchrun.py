"""The driver starts multi-GPU jobs with ``python -m torch.distributed.run``: the elastic agent then owns
MASTER_PORT, and this package's rendezvous must still come up (reference launch path: README.md:341-343)."""
import os
import subprocess
import sys

from conftest import free_port

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_env_rendezvous_under_torchrun(tmp_path):
    script = tmp_path / "w.py"
    script.write_text(
        "import os, sys, torch\n"
        "sys.path.insert(0, %r)\n"
        "import pytorch_distributed_learning_b200 as pdl\n"
        "pdl.distributed.init_process_group(backend='gloo')\n"
        "t = torch.ones(4) * (pdl.distributed.get_rank() + 1)\n"
        "pdl.distributed.all_reduce(t)\n"
        "open(os.path.join(%r, 'r' + os.environ['RANK']), 'w').write(str(t.tolist()))\n"
        "pdl.distributed.destroy_process_group()\n" % (ROOT, str(tmp_path)))
    rc = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                         "--master-port", str(free_port()), str(script)], cwd=ROOT, timeout=240).returncode
    assert rc == 0
    assert (tmp_path / "r0").read_text() == "[3.0, 3.0, 3.0, 3.0]" and (tmp_path / "r1").read_text() == "[3.0, 3.0, 3.0, 3.0]"
