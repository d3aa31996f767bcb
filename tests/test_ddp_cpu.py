"""BASELINE config 1: ConvNet, spawn launcher, CPU backend, world_size 2, synthetic 1x28x28."""
import pytest
import torch

import _workers
import pytorch_distributed_learning_b200 as pdl
from conftest import free_port


@pytest.mark.parametrize("momentum", [False, True])
def test_convnet_matches_torch_ddp_gloo(tmp_path, momentum):
    steps, world = 4, 2
    pdl.multiprocessing.spawn(_workers.train_ours, args=(world, free_port(), str(tmp_path), steps, momentum), nprocs=world)
    torch.multiprocessing.spawn(_workers.train_torch, args=(world, free_port(), str(tmp_path), steps, momentum), nprocs=world)
    ours = [torch.load(tmp_path / ("ours_r%d.pt" % r), weights_only=False) for r in range(world)]
    ref = [torch.load(tmp_path / ("torch_r%d.pt" % r), weights_only=False) for r in range(world)]
    # replicas must be bit-identical to each other
    assert torch.equal(ours[0]["params"], ours[1]["params"])
    # and follow the same trajectory as the reference's stack
    for r in range(world):
        torch.testing.assert_close(torch.tensor(ours[r]["losses"]), torch.tensor(ref[r]["losses"]), rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(ours[0]["params"], ref[0]["params"], rtol=1e-4, atol=1e-6)


def test_no_sync(tmp_path):
    pdl.multiprocessing.spawn(_workers.no_sync_worker, args=(2, free_port(), str(tmp_path)), nprocs=2)
    r0 = torch.load(tmp_path / "r0.pt", weights_only=False)
    r1 = torch.load(tmp_path / "r1.pt", weights_only=False)
    # inside no_sync the gradient is the local one: d(sum(Wx+b))/dW = sum over batch of x
    assert torch.equal(r0["local"], torch.full((3, 4), 2.0)) and torch.equal(r1["local"], torch.full((3, 4), 4.0))
    # the synchronising backward averages (accumulated local + new) over ranks
    expect = ((2.0 + 20.0) + (4.0 + 40.0)) / 2
    assert torch.equal(r0["synced"], torch.full((3, 4), expect)) and torch.equal(r1["synced"], r0["synced"])


def test_checkpoint_roundtrip(tmp_path):
    pdl.multiprocessing.spawn(_workers.checkpoint_worker, args=(2, free_port(), str(tmp_path)), nprocs=2)
    r0 = torch.load(tmp_path / "r0.pt", weights_only=False)
    r1 = torch.load(tmp_path / "r1.pt", weights_only=False)
    assert r0["wrote"] and not r1["wrote"]
    assert r0["roundtrip"] and r1["roundtrip"] and r0["epoch"] == 3
    # the plain file is a bare state_dict with the unwrapped module's keys, loadable by stock torch
    sd = torch.load(tmp_path / "ck.pt")
    assert list(sd.keys())[0] == "conv1.weight" and len(sd) == 8
    from pytorch_distributed_learning_b200.models import ConvNet

    ConvNet().load_state_dict(sd)


def test_ddp_argument_validation(port):
    dist = pdl.distributed
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, world_size=1, rank=0)
    try:
        frozen = torch.nn.Linear(2, 2)
        for p in frozen.parameters():
            p.requires_grad_(False)
        with pytest.raises(RuntimeError, match="not needed"):
            pdl.DistributedDataParallel(frozen)
        with pytest.raises(ValueError, match="device_ids"):
            pdl.DistributedDataParallel(torch.nn.Linear(2, 2), device_ids=[0])
        ddp = pdl.DistributedDataParallel(torch.nn.Linear(2, 2))
        assert ddp.fused_optimizer  # default on (only GPU buckets of optim.SGD can take it)
        assert not pdl.DistributedDataParallel(torch.nn.Linear(2, 2), fused_optimizer=False).fused_optimizer
        assert list(ddp.state_dict().keys()) == ["module.weight", "module.bias"]
        ddp(torch.ones(1, 2)).sum().backward()
        assert ddp.get_step_stats()["iterations"] == 1
    finally:
        dist.destroy_process_group()


def test_register_comm_hook(tmp_path):
    pdl.multiprocessing.spawn(_workers.comm_hook_worker, args=(2, free_port(), str(tmp_path)), nprocs=2)
    r0 = torch.load(tmp_path / "r0.pt", weights_only=False)
    r1 = torch.load(tmp_path / "r1.pt", weights_only=False)
    assert r0["twice_rejected"]
    for (plain0, hooked0), (plain1, hooked1) in zip(r0["grads"], r1["grads"]):
        torch.testing.assert_close(hooked0, plain0)
        assert torch.equal(hooked0, hooked1)  # replicas agree
    # tiny caps => several buckets per iteration; the hook saw each of them, the last one flagged
    per_iter = len(r0["seen"]) // 2
    assert per_iter >= 2 and r0["seen"][per_iter - 1][1] and not r0["seen"][0][1]


def test_unused_parameter_is_left_alone_by_the_optimizer(port):
    """A parameter that took no part in backward keeps ``grad is None`` and is not decayed (torch skips it)."""
    dist = pdl.distributed
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, world_size=1, rank=0)
    try:
        class Two(torch.nn.Module):
            def __init__(self):
                super().__init__()
                self.used, self.unused = torch.nn.Linear(3, 3), torch.nn.Linear(3, 3)

            def forward(self, x):
                return self.used(x)

        m = Two()
        before = m.unused.weight.detach().clone()
        opt = pdl.optim.SGD(m.parameters(), lr=0.1, weight_decay=0.5)
        ddp = pdl.DistributedDataParallel(m, find_unused_parameters=True)
        ddp(torch.ones(2, 3)).sum().backward()
        opt.step()
        assert m.unused.weight.grad is None and torch.equal(m.unused.weight.detach(), before)
        assert m.used.weight.grad is not None
    finally:
        dist.destroy_process_group()
