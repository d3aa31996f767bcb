"""Failure detection of the device-side collectives: a rank that never shows up must make its peers' kernels
give up after PDL_KERNEL_TIMEOUT_S and the next host synchronisation point raise -- not wedge the GPU."""
import os
import time

import pytest
import torch

pytestmark = pytest.mark.gpu

import pytorch_distributed_learning_b200 as pdl  # noqa: E402
from conftest import free_port  # noqa: E402

NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0


def _absent_peer_worker(rank, world, port, outdir):
    os.environ["PDL_KERNEL_TIMEOUT_S"] = "2"
    dist = pdl.distributed
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, world_size=world, rank=rank)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    pg = dist._default()
    t = torch.ones(4096, device=dev)
    dist.all_reduce(t)  # healthy collective first: maps the segments, both ranks take part
    torch.cuda.synchronize()
    res = {"healthy": bool((t == world).all().item()), "clean_before": pg.check_device_errors(raise_on_error=False) == 0}
    if rank == 0:
        # rank 1 never launches this one: the kernel must come back by itself
        t0 = time.time()
        dist.all_reduce(torch.ones(4096, device=dev))
        torch.cuda.synchronize()
        res["seconds"] = time.time() - t0
        try:
            pg.check_device_errors()
            res["raised"] = None
        except RuntimeError as e:
            res["raised"] = str(e)
    torch.save(res, os.path.join(outdir, "r%d.pt" % rank))
    # the group is inconsistent now (epochs differ): tear down without further collectives
    try:
        dist.destroy_process_group()
    except Exception:
        pass


@pytest.mark.skipif(NGPU < 2, reason="needs at least 2 GPUs")
def test_absent_peer_times_out_and_is_reported(tmp_path):
    pdl.multiprocessing.spawn(_absent_peer_worker, args=(2, free_port(), str(tmp_path)), nprocs=2)
    r0 = torch.load(tmp_path / "r0.pt", weights_only=False)
    r1 = torch.load(tmp_path / "r1.pt", weights_only=False)
    assert r0["healthy"] and r1["healthy"] and r0["clean_before"] and r1["clean_before"]
    assert 1.0 < r0["seconds"] < 30.0, "the kernel should give up after ~PDL_KERNEL_TIMEOUT_S, took %.1f s" % r0["seconds"]
    assert r0["raised"] and "timed out waiting for a peer" in r0["raised"]


def _lost_peer_training_worker(rank, world, port, outdir):
    os.environ["PDL_KERNEL_TIMEOUT_S"] = "2"
    dist = pdl.distributed
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, world_size=world, rank=rank)
    torch.cuda.set_device(rank)
    dev = torch.device("cuda", rank)
    torch.manual_seed(0)
    model = pdl.models.ConvNet().to(dev)
    opt = pdl.optim.SGD(model.parameters(), lr=0.1, momentum=0.9)
    ddp = pdl.nn.parallel.DistributedDataParallel(model, device_ids=[rank])
    step = pdl.TrainStep(ddp, pdl.nn.CrossEntropyLoss(), opt)
    x, y = torch.randn(16, 1, 28, 28, device=dev), torch.randint(0, 10, (16,), device=dev)
    for _ in range(2):
        step(x, y)  # healthy steps: both ranks
    torch.cuda.synchronize()
    res = {}
    if rank == 0:
        before = [p.detach().clone() for p in model.parameters()]
        mom_before = [opt.state[p]["momentum_buffer"].detach().clone() for p in model.parameters()]
        step(x, y)  # rank 1 never replays this one: the fused allreduce+SGD kernel must give up WITHOUT touching the weights
        torch.cuda.synchronize()
        res["params_unchanged"] = all(torch.equal(a, b.detach()) for a, b in zip(before, model.parameters()))
        res["momentum_unchanged"] = all(torch.equal(a, opt.state[p]["momentum_buffer"]) for a, p in zip(mom_before, model.parameters()))
        try:
            step(x, y)  # the step path polls the pinned error word: raises without any synchronisation
            res["raised"] = None
        except RuntimeError as e:
            res["raised"] = str(e)
        torch.cuda.synchronize()
        res["still_unchanged"] = all(torch.equal(a, b.detach()) for a, b in zip(before, model.parameters()))
    torch.save(res, os.path.join(outdir, "r%d.pt" % rank))
    os._exit(0)  # the group is inconsistent: no collective teardown


@pytest.mark.skipif(NGPU < 2, reason="needs at least 2 GPUs")
def test_lost_peer_leaves_parameters_untouched(tmp_path):
    pdl.multiprocessing.spawn(_lost_peer_training_worker, args=(2, free_port(), str(tmp_path)), nprocs=2)
    r0 = torch.load(tmp_path / "r0.pt", weights_only=False)
    assert r0["params_unchanged"] and r0["momentum_unchanged"], r0
    assert r0["raised"] and "timed out waiting for a peer" in r0["raised"], r0
    assert r0["still_unchanged"], r0
