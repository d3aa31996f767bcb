"""Multi-GPU checks (>= 2 B200s on one box): peer-memory collectives and DDP end to end."""
import copy
import os

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

import pytorch_distributed_learning_b200 as pdl  # noqa: E402
from conftest import free_port  # noqa: E402

NGPU = torch.cuda.device_count() if torch.cuda.is_available() else 0
needs2 = pytest.mark.skipif(NGPU < 2, reason="needs at least 2 GPUs")


def _collectives_worker(rank, world, port, outdir):
    dist = pdl.distributed
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % port, world_size=world, rank=rank)
    torch.cuda.set_device(rank)  # AFTER init, like mpspawn_dist.py:49-60
    dev = torch.device("cuda", rank)
    pg = dist._default()
    res = {}
    for name in ["oneshot", "twoshot", "ring", "push", "nvls"]:
        for n in (7, 1000, 113674, 3_000_001):
            g = torch.Generator().manual_seed(100 + n)
            parts = [torch.randn(n, generator=g) for _ in range(world)]
            t = parts[rank].to(dev)
            try:
                pg.allreduce(t, dist.ReduceOp.SUM, algorithm=name)
            except RuntimeError as e:
                if name == "nvls" and "multicast" in str(e):
                    res["nvls"] = "unavailable"
                    break
                raise
            expect = torch.stack(parts).double().sum(0)
            err = (t.cpu().double() - expect).abs().max().item()
            res["%s_%d" % (name, n)] = err
            everyone = pg.allgather(t)
            res["%s_%d_identical" % (name, n)] = all(torch.equal(everyone[0], e) for e in everyone)
    b = torch.full((100003,), float(rank), device=dev)
    dist.broadcast(b, src=world - 1)
    res["bcast"] = bool((b == world - 1).all().item())
    parts = [torch.zeros(5, device=dev) for _ in range(world)]
    dist.all_gather(parts, torch.full((5,), float(rank), device=dev))
    res["gather"] = [p[0].item() for p in parts]
    m = torch.tensor([rank], device=dev)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    res["max"] = m.item()
    for _ in range(200):  # flag/epoch soak
        s = torch.ones(1024, device=dev)
        dist.all_reduce(s)
    res["soak"] = bool((s == world).all().item())
    torch.save(res, os.path.join(outdir, "r%d.pt" % rank))
    dist.destroy_process_group()


@needs2
@pytest.mark.parametrize("world", [2] + ([NGPU] if NGPU > 2 else []))
def test_peer_memory_collectives(tmp_path, world):
    pdl.multiprocessing.spawn(_collectives_worker, args=(world, free_port(), str(tmp_path)), nprocs=world)
    for r in range(world):
        res = torch.load(tmp_path / ("r%d.pt" % r), weights_only=False)
        for k, v in res.items():
            if k.endswith("_identical"):
                assert v, k
            elif k.split("_")[0] in ("oneshot", "twoshot", "ring", "push", "nvls") and isinstance(v, float):
                assert v < 1e-4, (k, v)
        assert res["bcast"] and res["soak"] and res["max"] == world - 1
        assert res["gather"] == [float(i) for i in range(world)]


class RefConvNet(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.conv1, self.conv2, self.conv3 = torch.nn.Conv2d(1, 32, 5, 1, 1), torch.nn.Conv2d(32, 64, 3), torch.nn.Conv2d(64, 128, 3)
        self.fc1 = torch.nn.Linear(2048, 10)

    def forward(self, x):
        x = F.max_pool2d(torch.relu(self.conv1(x)), 2, 2)
        x = F.max_pool2d(torch.relu(self.conv2(x)), 2, 1)
        x = F.max_pool2d(torch.relu(self.conv3(x)), 2, 2)
        return self.fc1(x.view(-1, 2048))


def _ddp_worker(rank, world, port, outdir, mode, algorithm):
    os.environ.update({"MASTER_ADDR": "127.0.0.1", "MASTER_PORT": str(port), "RANK": str(rank), "WORLD_SIZE": str(world), "LOCAL_RANK": str(rank)})
    tf32 = mode.startswith("graph_tf32")  # the tensor-core schedule (four-part gradient exchange); everything else runs strict fp32
    if mode == "graph_tf32_fusec2":
        os.environ["PDL_CONVNET_FUSE_C2"] = "1"  # conv2's exchange inside the second pass of its weight gradient
    torch.backends.cudnn.allow_tf32 = tf32
    torch.backends.cuda.matmul.allow_tf32 = False
    dist = pdl.distributed
    torch.cuda.set_device(rank % torch.cuda.device_count())  # BEFORE init, like launch_dist.py:47-49
    dist.init_process_group(backend="nccl")
    torch.manual_seed(10 + rank)  # different on purpose: rank 0's weights must win
    model = pdl.models.ConvNet().to(torch.device("cuda", rank))
    kw = dict(lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)
    opt = pdl.optim.SGD(model.parameters(), **kw) if mode != "torch_optim" else torch.optim.SGD(model.parameters(), **kw)
    ddp = pdl.nn.parallel.DistributedDataParallel(model, device_ids=[rank], output_device=rank, algorithm=algorithm,
                                                  graph_step=None if mode == "sixcall" else False)  # "eager": the call-by-call hook path
    crit = pdl.nn.CrossEntropyLoss().cuda(rank)
    start = copy.deepcopy({k: v.cpu() for k, v in model.state_dict().items()})
    step = pdl.TrainStep(ddp, crit, opt) if mode.startswith("graph") else None
    g = torch.Generator().manual_seed(55)
    losses = []
    for _ in range(4):
        x = torch.randn(world, 10, 1, 28, 28, generator=g)
        y = torch.randint(0, 10, (world, 10), generator=g)
        if step is not None:
            losses.append(step(x[rank].pin_memory(), y[rank].pin_memory()).item())
        else:
            loss = crit(ddp(x[rank].cuda(rank, non_blocking=True)), y[rank].cuda(rank, non_blocking=True))
            opt.zero_grad()
            loss.backward()
            opt.step()
            losses.append(loss.item())
    torch.cuda.synchronize()
    torch.save({"start": start, "end": {k: v.cpu() for k, v in model.state_dict().items()}, "losses": losses, "stats": ddp.get_step_stats()},
               os.path.join(outdir, "r%d.pt" % rank))
    dist.destroy_process_group()


@needs2
@pytest.mark.parametrize("mode,algorithm", [("graph", None), ("graph_tf32", None), ("graph_tf32_fusec2", None), ("sixcall", None), ("eager", None), ("eager", "twoshot"), ("eager", "ring"),
                                            ("eager", "push"), ("eager", "oneshot"), ("torch_optim", None)])
def test_ddp_convnet_two_gpus(tmp_path, mode, algorithm):
    world = 2
    pdl.multiprocessing.spawn(_ddp_worker, args=(world, free_port(), str(tmp_path), mode, algorithm), nprocs=world)
    res = [torch.load(tmp_path / ("r%d.pt" % r), weights_only=False) for r in range(world)]
    for k in res[0]["end"]:
        assert torch.equal(res[0]["start"][k], res[1]["start"][k]), "constructor broadcast must equalise replicas"
        assert torch.equal(res[0]["end"][k], res[1]["end"][k]), "replicas drifted: %s" % k
    # single-process reference on the concatenated batch
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    ref = RefConvNet().cuda()
    ref.load_state_dict(res[0]["start"])
    opt = torch.optim.SGD(ref.parameters(), lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)
    g = torch.Generator().manual_seed(55)
    for _ in range(4):
        x = torch.randn(world, 10, 1, 28, 28, generator=g).flatten(0, 1).cuda()
        y = torch.randint(0, 10, (world, 10), generator=g).flatten(0, 1).cuda()
        opt.zero_grad()
        F.cross_entropy(ref(x), y).backward()
        opt.step()
    tf32 = mode.startswith("graph_tf32")  # TF32 convolutions against the strict-fp32 oracle: TF32 tolerance
    for k, v in ref.state_dict().items():
        torch.testing.assert_close(res[0]["end"][k], v.cpu(), rtol=2e-2 if tf32 else 1e-4, atol=2e-3 if tf32 else 1e-5, msg=lambda m, k=k: "%s: %s" % (k, m))
    if mode != "torch_optim":
        assert res[0]["stats"].get("fused_buckets", 0) > 0 or mode.startswith("graph") or mode == "sixcall"
