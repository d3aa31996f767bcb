import pytest
import torch

import _workers
import pytorch_distributed_learning_b200 as pdl
from conftest import free_port


@pytest.mark.parametrize("world,method", [(2, "env"), (3, "tcp"), (4, "env")])
def test_collectives(tmp_path, world, method):
    pdl.multiprocessing.spawn(_workers.collectives, args=(world, free_port(), method, str(tmp_path)), nprocs=world)
    for r in range(world):
        res = torch.load(tmp_path / ("r%d.pt" % r), weights_only=False)
        assert torch.equal(res["sum"], torch.arange(10, dtype=torch.float32) * world + sum(range(world)))
        assert res["big_ok"]
        assert res["max"].tolist() == [world - 1, 0]
        assert res["bcast"].tolist() == [float(world - 1)] * 5
        assert res["gather"].tolist() == [[i] * 3 for i in range(world)]
        assert res["objs"] == [{"rank": i, "s": "x" * (i + 1)} for i in range(world)]
        assert res["bol"] == [0, "hello"]
        assert (res["rank"], res["world"], res["backend"]) == (r, world, "gloo")


def test_new_group(tmp_path):
    pdl.multiprocessing.spawn(_workers.subgroups, args=(4, free_port(), str(tmp_path)), nprocs=4)
    for r in range(4):
        res = torch.load(tmp_path / ("r%d.pt" % r), weights_only=False)
        assert res["sub"].item() == (2.0 if r % 2 == 0 else 4.0)
        assert res["world"].item() == 4.0
        assert res["gsize"] == 2 and res["grank"] == r // 2
        assert res["ranks"] == ([0, 2] if r % 2 == 0 else [1, 3])


def test_double_init_and_uninitialised(port, monkeypatch):
    dist = pdl.distributed
    with pytest.raises(ValueError, match="not been initialized"):
        dist.get_rank()
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, world_size=1, rank=0)
    try:
        with pytest.raises(ValueError, match="twice"):
            dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, world_size=1, rank=0)
        assert dist.get_world_size() == 1 and dist.get_rank() == 0 and dist.is_initialized()
        t = torch.ones(3)
        dist.all_reduce(t)
        assert t.tolist() == [1, 1, 1]
    finally:
        dist.destroy_process_group()
    assert not dist.is_initialized()
    with pytest.raises(RuntimeError, match="MPI"):
        dist.init_process_group("mpi", init_method="tcp://127.0.0.1:%d" % port, world_size=1, rank=0)
    with pytest.raises(ValueError, match="both"):
        dist.init_process_group("gloo", init_method="env://", store=object(), world_size=1, rank=0)


def test_gather_scatter_reduce_scatter(tmp_path):
    world = 3
    pdl.multiprocessing.spawn(_workers.more_collectives, args=(world, free_port(), str(tmp_path)), nprocs=world)
    for r in range(world):
        res = torch.load(tmp_path / ("r%d.pt" % r), weights_only=False)
        assert res["agit"].tolist() == [[float(i)] * 3 for i in range(world)]
        full = torch.arange(2 * world, dtype=torch.float32) * world + sum(range(world))
        assert torch.equal(res["rs"], full.view(world, 2)[r])
        assert (res["gather"] is None) == (r != 1)
        if r == 1:
            assert res["gather"].tolist() == [[float(i + 1)] * 2 for i in range(world)]
        assert res["scatter"].tolist() == [10.0 * r] * 2


@pytest.mark.parametrize("mode", ["stuck", "dead"])
def test_collective_fails_instead_of_hanging(tmp_path, mode):
    """A peer that is alive but never joins is reported after the group's timeout (3 s here); a peer that exited is
    reported at once.  (torch gets the former from its watchdog, the latter from the broken connection.)"""
    pdl.multiprocessing.spawn(_workers.stuck_peer, args=(2, free_port(), str(tmp_path), mode), nprocs=2)
    waited, msg = (tmp_path / "r0.txt").read_text().split("|", 1)
    if mode == "stuck":
        assert "timed out after 3 s" in msg and 2.5 <= float(waited) < 9.0, (waited, msg)
    else:
        assert "a peer exited" in msg and float(waited) < 3.0, (waited, msg)
