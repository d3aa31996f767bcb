"""Worker functions for the multi-process tests (must be importable by spawned children)."""
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import pytorch_distributed_learning_b200 as pdl  # noqa: E402
from pytorch_distributed_learning_b200.models import ConvNet  # noqa: E402

dist = pdl.distributed


def _init(rank, world, port, method="env"):
    if method == "env":
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", init_method="env://", world_size=world, rank=rank)
    else:
        dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, world_size=world, rank=rank)


def stuck_peer(rank, world, port, outdir, mode):
    """Rank 1 never joins the second allreduce: it sleeps ("stuck") or exits ("dead").  Rank 0 must get an error, not hang."""
    import time
    from datetime import timedelta

    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", init_method="env://", world_size=world, rank=rank, timeout=timedelta(seconds=3))
    t = torch.ones(4)
    dist.all_reduce(t)  # everyone takes part in the first one
    if rank == 1:
        if mode == "stuck":
            time.sleep(8)
        os._exit(0)
    t0 = time.time()
    try:
        dist.all_reduce(t)
        msg = "no error"
    except RuntimeError as e:
        msg = str(e)
    with open(os.path.join(outdir, "r0.txt"), "w") as f:
        f.write("%.1f|%s" % (time.time() - t0, msg))
    os._exit(0)


def collectives(rank, world, port, method, outdir):
    _init(rank, world, port, method)
    res = {}
    t = torch.arange(10, dtype=torch.float32) + rank
    dist.all_reduce(t)
    res["sum"] = t.clone()
    big = torch.full((100003,), float(rank + 1), dtype=torch.float64)
    dist.all_reduce(big)
    res["big_ok"] = bool((big == sum(range(1, world + 1))).all())
    m = torch.tensor([rank, -rank], dtype=torch.int64)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    res["max"] = m.clone()
    b = torch.full((5,), float(rank))
    dist.broadcast(b, src=world - 1)
    res["bcast"] = b.clone()
    parts = [torch.zeros(3, dtype=torch.int32) for _ in range(world)]
    dist.all_gather(parts, torch.full((3,), rank, dtype=torch.int32))
    res["gather"] = torch.stack(parts)
    objs = [None] * world
    dist.all_gather_object(objs, {"rank": rank, "s": "x" * (rank + 1)})
    res["objs"] = objs
    lst = [rank, "hello"] if rank == 0 else [None, None]
    dist.broadcast_object_list(lst, src=0)
    res["bol"] = lst
    dist.barrier()
    res["rank"], res["world"], res["backend"] = dist.get_rank(), dist.get_world_size(), dist.get_backend()
    torch.save(res, os.path.join(outdir, "r%d.pt" % rank))
    dist.destroy_process_group()


def subgroups(rank, world, port, outdir):
    _init(rank, world, port)
    g_even = dist.new_group([0, 2])
    g_odd = dist.new_group([1, 3])
    mine = g_even if rank % 2 == 0 else g_odd
    other = g_odd if rank % 2 == 0 else g_even
    assert other is dist.GroupMember.NON_GROUP_MEMBER
    t = torch.tensor([float(rank)])
    dist.all_reduce(t, group=mine)
    w = torch.tensor([1.0])
    dist.all_reduce(w)
    torch.save({"sub": t, "world": w, "grank": dist.get_rank(mine), "gsize": dist.get_world_size(mine),
                "ranks": dist.get_process_group_ranks(mine)}, os.path.join(outdir, "r%d.pt" % rank))
    dist.destroy_process_group()


def _train(rank, world, port, outdir, steps, use_ours, momentum):
    torch.manual_seed(1234 + rank)  # deliberately different: the constructor broadcast must equalise
    torch.set_num_threads(2)
    model = ConvNet()
    if use_ours:
        _init(rank, world, port)
        kw = dict(momentum=0.9, weight_decay=1e-4, nesterov=True) if momentum else {}
        opt = pdl.optim.SGD(model.parameters(), 1e-2, **kw)
        ddp = pdl.nn.parallel.DistributedDataParallel(model)
        crit = pdl.nn.CrossEntropyLoss()
    else:
        import torch.distributed as tdist

        tdist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, world_size=world, rank=rank)
        kw = dict(momentum=0.9, weight_decay=1e-4, nesterov=True) if momentum else {}
        opt = torch.optim.SGD(model.parameters(), 1e-2, **kw)
        ddp = torch.nn.parallel.DistributedDataParallel(model)
        crit = torch.nn.CrossEntropyLoss()
    if rank == 0 and use_ours:
        pass
    # both implementations must start from rank 0's weights: seed 1234
    g = torch.Generator().manual_seed(77 + rank)
    losses = []
    for _ in range(steps):
        x = torch.randn(16, 1, 28, 28, generator=g)
        y = torch.randint(0, 10, (16,), generator=g)
        loss = crit(ddp(x), y)
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(float(loss.detach()))
    flat = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    torch.save({"params": flat, "losses": losses}, os.path.join(outdir, "%s_r%d.pt" % ("ours" if use_ours else "torch", rank)))
    if use_ours:
        dist.destroy_process_group()
    else:
        tdist.destroy_process_group()


def train_ours(rank, world, port, outdir, steps, momentum):
    _train(rank, world, port, outdir, steps, True, momentum)


def train_torch(rank, world, port, outdir, steps, momentum):
    _train(rank, world, port, outdir, steps, False, momentum)


def failing(rank, flag_dir):
    import time

    if rank == 1:
        raise ValueError("boom from rank 1")
    time.sleep(60)
    open(os.path.join(flag_dir, "survivor"), "w").close()


def exiting(rank):
    if rank == 0:
        os._exit(7)
    import time

    time.sleep(60)


def ok_worker(rank, outdir):
    open(os.path.join(outdir, "ok%d" % rank), "w").close()


def no_sync_worker(rank, world, port, outdir):
    _init(rank, world, port)
    torch.manual_seed(0)
    model = torch.nn.Linear(4, 3)
    ddp = pdl.DistributedDataParallel(model)
    x1 = torch.full((2, 4), float(rank + 1))
    x2 = torch.full((2, 4), float(10 * (rank + 1)))
    with ddp.no_sync():
        ddp(x1).sum().backward()
    local_only = model.weight.grad.clone()
    ddp(x2).sum().backward()
    torch.save({"local": local_only, "synced": model.weight.grad.clone()}, os.path.join(outdir, "r%d.pt" % rank))
    dist.destroy_process_group()


def checkpoint_worker(rank, world, port, outdir):
    _init(rank, world, port)
    from pytorch_distributed_learning_b200.utils.checkpoint import load_checkpoint, save_checkpoint

    torch.manual_seed(rank)
    model = ConvNet()
    ddp = pdl.DistributedDataParallel(model)
    opt = pdl.optim.SGD(model.parameters(), 0.1, momentum=0.9)
    crit = pdl.nn.CrossEntropyLoss()
    x, y = torch.randn(4, 1, 28, 28), torch.randint(0, 10, (4,))
    crit(ddp(x), y).backward()
    opt.step()
    path = os.path.join(outdir, "ck.pt")
    wrote = save_checkpoint(ddp, path)
    full = os.path.join(outdir, "full.pt")
    save_checkpoint(ddp, full, optimizer=opt, epoch=3)
    before = torch.cat([p.detach().reshape(-1) for p in model.parameters()]).clone()
    with torch.no_grad():
        for p in model.parameters():
            p.add_(1.0)
    info = load_checkpoint(ddp, full, optimizer=opt)
    after = torch.cat([p.detach().reshape(-1) for p in model.parameters()])
    torch.save({"wrote": wrote, "roundtrip": torch.equal(before, after), "epoch": info.get("epoch")}, os.path.join(outdir, "r%d.pt" % rank))
    dist.destroy_process_group()


class _Later:
    """Minimal future: what a hook may return instead of a tensor."""

    def __init__(self, fn):
        self.fn = fn

    def wait(self):
        return self.fn()


def comm_hook_worker(rank, world, port, outdir):
    """DDP with a user communication hook (allreduce the un-divided bucket, scale by 1/W, hand back a future) must
    produce the same gradients as the built-in reducer."""
    _init(rank, world, port)
    torch.manual_seed(0)
    a = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    b = torch.nn.Sequential(torch.nn.Linear(6, 5), torch.nn.Tanh(), torch.nn.Linear(5, 3))
    b.load_state_dict(a.state_dict())
    plain, hooked = pdl.DistributedDataParallel(a), pdl.DistributedDataParallel(b, bucket_cap_mb=1e-5, first_bucket_mb=1e-5)
    seen = []

    def hook(state, bucket):
        seen.append((bucket.index(), bucket.is_last(), bucket.buffer().numel(), len(bucket.gradients())))
        buf = bucket.buffer()

        def finish():
            dist.all_reduce(buf)
            return buf.div_(state["world"])
        return _Later(finish)

    hooked.register_comm_hook({"world": world}, hook)
    try:
        hooked.register_comm_hook(None, hook)
        twice = False
    except RuntimeError:
        twice = True
    grads = []
    for it in range(2):
        x = torch.randn(4, 6, generator=torch.Generator().manual_seed(10 * it + rank))
        for m in (plain, hooked):
            m.zero_grad()
            m(x).pow(2).sum().backward()
        grads.append((torch.cat([p.grad.reshape(-1) for p in a.parameters()]), torch.cat([p.grad.reshape(-1) for p in b.parameters()])))
    torch.save({"grads": grads, "seen": seen, "twice_rejected": twice}, os.path.join(outdir, "r%d.pt" % rank))
    dist.destroy_process_group()


def tracing_worker(rank, world, port, outdir):
    os.environ["PDL_TRACE"] = os.path.join(outdir, "trace")
    os.environ["RANK"] = str(rank)
    from pytorch_distributed_learning_b200.utils import tracing

    tracing.reset()
    _init(rank, world, port)
    torch.manual_seed(0)
    ddp = pdl.DistributedDataParallel(ConvNet())
    step = pdl.TrainStep(ddp, pdl.nn.CrossEntropyLoss(), pdl.optim.SGD(ddp.parameters(), lr=0.01))
    for _ in range(2):
        step(torch.randn(4, 1, 28, 28), torch.randint(0, 10, (4,)))
    tracing.flush()
    dist.destroy_process_group()


def more_collectives(rank, world, port, outdir):
    _init(rank, world, port)
    res = {}
    out = torch.zeros(world, 3)
    dist.all_gather_into_tensor(out, torch.full((3,), float(rank)))
    res["agit"] = out.clone()
    rs = torch.zeros(2)
    dist.reduce_scatter_tensor(rs, torch.arange(2 * world, dtype=torch.float32) + rank)
    res["rs"] = rs.clone()
    glist = [torch.zeros(2) for _ in range(world)] if rank == 1 else None
    dist.gather(torch.full((2,), float(rank + 1)), glist, dst=1)
    res["gather"] = torch.stack(glist) if glist is not None else None
    mine = torch.zeros(2)
    dist.scatter(mine, [torch.full((2,), 10.0 * i) for i in range(world)] if rank == 0 else None, src=0)
    res["scatter"] = mine.clone()
    torch.save(res, os.path.join(outdir, "r%d.pt" % rank))
    dist.destroy_process_group()
