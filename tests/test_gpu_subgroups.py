"""new_group on the NVLink backend: a sub-communicator owns its symmetric arena and signal channel, collectives on it do
not disturb the default group, and non-members get the sentinel (reference README.md:28,39 names new_group as surface)."""
import os

import pytest
import torch

pytestmark = pytest.mark.gpu

import pytorch_distributed_learning_b200 as pdl  # noqa: E402
from conftest import free_port  # noqa: E402

NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0


def _worker(rank, world, port, outdir):
    dist = pdl.distributed
    torch.cuda.set_device(rank)
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, world_size=world, rank=rank)
    dev = torch.device("cuda", rank)
    res = {}
    both = dist.new_group([0, 1])  # a second communicator over the same ranks
    solo = [dist.new_group([r]) for r in range(world)]  # every rank calls every new_group, members or not
    res["sentinel"] = all((g is dist.GroupMember.NON_GROUP_MEMBER) == (i != rank) for i, g in enumerate(solo))
    a = torch.full((1000,), float(rank + 1), device=dev)
    dist.all_reduce(a, group=both)
    res["both"] = a.cpu()
    b = torch.full((10,), float(rank + 5), device=dev)
    dist.all_reduce(b, group=solo[rank])  # a group of one: identity
    res["solo"] = b.cpu()
    res["sizes"] = (dist.get_world_size(both), dist.get_rank(both), dist.get_world_size(solo[rank]), dist.get_rank(solo[rank]))
    c = torch.arange(6, dtype=torch.float32, device=dev) * (rank + 1)
    dist.all_reduce(c)  # the default group still works, interleaved with the sub-group
    dist.all_reduce(a, group=both)
    res["default"], res["both2"] = c.cpu(), a.cpu()
    bc = torch.full((4,), float(rank), device=dev)
    dist.broadcast(bc, src=1, group=both)
    res["bcast"] = bc.cpu()
    dist.destroy_process_group(both)
    dist.barrier()
    torch.save(res, os.path.join(outdir, "r%d.pt" % rank))
    dist.destroy_process_group()


@pytest.mark.skipif(NGPU < 2, reason="needs at least 2 GPUs")
def test_new_group_on_two_gpus(tmp_path):
    pdl.multiprocessing.spawn(_worker, args=(2, free_port(), str(tmp_path)), nprocs=2)
    for r in range(2):
        res = torch.load(tmp_path / ("r%d.pt" % r), weights_only=False)
        assert res["sentinel"]
        assert torch.equal(res["both"], torch.full((1000,), 3.0)) and torch.equal(res["both2"], torch.full((1000,), 6.0))
        assert torch.equal(res["solo"], torch.full((10,), float(r + 5)))
        assert res["sizes"] == (2, r, 1, 0)
        assert torch.equal(res["default"], torch.arange(6, dtype=torch.float32) * 3)
        assert torch.equal(res["bcast"], torch.full((4,), 1.0))
