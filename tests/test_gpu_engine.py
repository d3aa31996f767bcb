"""Whole-step checks on one B200: graphed TrainStep and the GPU DDP path (world size 1)."""
import copy

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

import pytorch_distributed_learning_b200 as pdl  # noqa: E402
from conftest import free_port  # noqa: E402


@pytest.fixture(autouse=True)
def _strict_fp32():
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


class RefConvNet(torch.nn.Module):  # the reference model's layer list (launch_dist.py:9-41) on stock torch ops
    def __init__(self):
        super().__init__()
        self.conv1, self.conv2, self.conv3 = torch.nn.Conv2d(1, 32, 5, 1, 1), torch.nn.Conv2d(32, 64, 3), torch.nn.Conv2d(64, 128, 3)
        self.fc1 = torch.nn.Linear(2048, 10)

    def forward(self, x):
        x = F.max_pool2d(torch.relu(self.conv1(x)), 2, 2)
        x = F.max_pool2d(torch.relu(self.conv2(x)), 2, 1)
        x = F.max_pool2d(torch.relu(self.conv3(x)), 2, 2)
        return self.fc1(x.view(-1, 2048))


def _reference_run(state, batches, opt_kw):
    ref = RefConvNet().cuda()
    ref.load_state_dict(state)
    opt = torch.optim.SGD(ref.parameters(), **opt_kw)
    losses = []
    for x, y in batches:
        loss = F.cross_entropy(ref(x.cuda()), y.cuda())
        opt.zero_grad()
        loss.backward()
        opt.step()
        losses.append(loss.item())
    return ref, losses


@pytest.mark.parametrize("opt_kw", [dict(lr=1e-2), dict(lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)])
@pytest.mark.parametrize("batch", [100, 7])
def test_graphed_trainstep_matches_torch(opt_kw, batch):
    torch.manual_seed(0)
    model = pdl.models.ConvNet().cuda()
    state = copy.deepcopy(model.state_dict())
    opt = pdl.optim.SGD(model.parameters(), **opt_kw)
    step = pdl.TrainStep(model, pdl.nn.CrossEntropyLoss(), opt)
    assert step.graphed
    g = torch.Generator().manual_seed(1)
    batches = [(torch.randn(batch, 1, 28, 28, generator=g).pin_memory(), torch.randint(0, 10, (batch,), generator=g).pin_memory()) for _ in range(5)]
    losses = [step(x, y).item() for x, y in batches]
    ref, ref_losses = _reference_run(state, batches, opt_kw)
    torch.testing.assert_close(torch.tensor(losses), torch.tensor(ref_losses), rtol=1e-4, atol=1e-5)
    # atol: two fp32 implementations can pick different arg-maxes for pool windows whose candidates differ by an ulp
    # (each implementation disagrees with an fp64 oracle about once per ~1e6 windows, scripts/dbg_strict.py); one such
    # flip moves a few weight gradients by ~5e-5, i.e. a parameter by lr/(1-momentum) times that
    for (n, a), (_, b) in zip(model.named_parameters(), ref.named_parameters()):
        torch.testing.assert_close(a, b, rtol=1e-4, atol=5e-5, msg=lambda m, n=n: "%s: %s" % (n, m))
    assert step.kernels_per_step() > 0
    # state_dict keys are the reference model's
    assert list(model.state_dict().keys())[0] == "conv1.weight"


def test_eager_trainstep_equals_graphed():
    torch.manual_seed(0)
    a = pdl.models.ConvNet().cuda()
    b = copy.deepcopy(a)
    oa, ob = pdl.optim.SGD(a.parameters(), 1e-2), pdl.optim.SGD(b.parameters(), 1e-2)
    sa = pdl.TrainStep(a, pdl.nn.CrossEntropyLoss(), oa)
    sb = pdl.TrainStep(b, pdl.nn.CrossEntropyLoss(), ob, use_graph=False)
    assert sa.graphed and not sb.graphed
    x, y = torch.randn(16, 1, 28, 28).cuda(), torch.randint(0, 10, (16,)).cuda()
    for _ in range(3):
        la, lb = sa(x, y), sb(x, y)
        torch.testing.assert_close(la, lb, rtol=1e-5, atol=1e-6)
    for pa, pb in zip(a.parameters(), b.parameters()):
        torch.testing.assert_close(pa, pb, rtol=1e-5, atol=1e-6)


@pytest.mark.parametrize("ours_optim", [True, False])
def test_ddp_world1_b200_backend(ours_optim):
    """Symmetric segment + bucket kernel on a single rank: fused SGD with our optimizer, plain
    averaging (then torch.optim.SGD) otherwise.  Reference call shape: launch_dist.py:59-61."""
    dist = pdl.distributed
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % free_port(), world_size=1, rank=0)
    try:
        torch.manual_seed(0)
        model = pdl.models.ConvNet().cuda()
        state = copy.deepcopy(model.state_dict())
        kw = dict(lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)
        opt = (pdl.optim.SGD if ours_optim else torch.optim.SGD)(model.parameters(), **kw)  # optimizer BEFORE the wrap, like launch_dist.py
        ddp = pdl.nn.parallel.DistributedDataParallel(model, device_ids=[0], output_device=0, graph_step=False)  # call by call: the hook path
        crit = pdl.nn.CrossEntropyLoss().cuda(0)
        g = torch.Generator().manual_seed(1)
        batches = [(torch.randn(12, 1, 28, 28, generator=g), torch.randint(0, 10, (12,), generator=g)) for _ in range(4)]
        losses = []
        for x, y in batches:
            x, y = x.cuda(non_blocking=True), y.cuda(non_blocking=True)
            loss = crit(ddp(x), y)
            opt.zero_grad()
            loss.backward()
            opt.step()
            losses.append(loss.item())
        stats = ddp.get_step_stats()
        assert stats["buckets_launched"] == 4 and stats["backend"] == "b200"
        assert stats["fused_buckets"] == (4 if ours_optim else 0)
        ref, ref_losses = _reference_run(state, batches, kw)
        torch.testing.assert_close(torch.tensor(losses), torch.tensor(ref_losses), rtol=1e-4, atol=1e-5)
        for a, b in zip(model.parameters(), ref.parameters()):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-5)
        # graphed step on top of the same DDP object keeps working
        if ours_optim:
            step = pdl.TrainStep(ddp, crit, opt)
            assert step.graphed
            l0 = step(batches[0][0].pin_memory(), batches[0][1].pin_memory()).item()
            assert l0 == l0
    finally:
        dist.destroy_process_group()


def test_generic_graph_engine_resnet18_matches_eager():
    """Any model: TrainStep captures the eager iteration (our ops, DDP hooks, fused SGD) into one CUDA graph."""
    dist = pdl.distributed
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % free_port(), world_size=1, rank=0)
    try:
        torch.manual_seed(0)
        kw = dict(lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)
        a = pdl.models.resnet18(num_classes=10).cuda()
        b = copy.deepcopy(a)
        da = pdl.nn.parallel.DistributedDataParallel(a, device_ids=[0])
        db = pdl.nn.parallel.DistributedDataParallel(b, device_ids=[0])
        oa, ob = pdl.optim.SGD(a.parameters(), **kw), pdl.optim.SGD(b.parameters(), **kw)
        crit = pdl.nn.CrossEntropyLoss()
        sa = pdl.TrainStep(da, crit, oa)
        sb = pdl.TrainStep(db, crit, ob, use_graph=False)
        assert sa.graphed and not sb.graphed
        g = torch.Generator().manual_seed(3)
        for _ in range(3):
            x, y = torch.randn(16, 3, 32, 32, generator=g).cuda(), torch.randint(0, 10, (16,), generator=g).cuda()
            la, lb = sa(x, y).item(), sb(x, y).item()
            assert abs(la - lb) <= 1e-3 * max(1.0, abs(lb)), (la, lb)
        assert sa.kernels_per_step() > 50
        for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
            torch.testing.assert_close(pa, pb, rtol=1e-3, atol=1e-4, msg=lambda m, n=n: "%s: %s" % (n, m))
        for (n, ba), (_, bb) in zip(a.named_buffers(), b.named_buffers()):
            torch.testing.assert_close(ba.float(), bb.float(), rtol=1e-3, atol=1e-4, msg=lambda m, n=n: "%s: %s" % (n, m))
    finally:
        dist.destroy_process_group()


def test_generic_graph_engine_without_ddp():
    """No DDP wrapper: gradients are fresh allocations and the stand-alone multi-tensor SGD builds its pointer table
    while the graph is being captured (pinned staging); replays must keep following the eager trajectory."""
    torch.manual_seed(0)
    kw = dict(lr=0.05, momentum=0.9, weight_decay=1e-4, nesterov=True)
    a = pdl.models.resnet18(num_classes=10).cuda()
    b = copy.deepcopy(a)
    oa, ob = pdl.optim.SGD(a.parameters(), **kw), pdl.optim.SGD(b.parameters(), **kw)
    crit = pdl.nn.CrossEntropyLoss()
    sa, sb = pdl.TrainStep(a, crit, oa), pdl.TrainStep(b, crit, ob, use_graph=False)
    assert sa.graphed and not sb.graphed
    g = torch.Generator().manual_seed(5)
    for _ in range(4):
        x, y = torch.randn(8, 3, 32, 32, generator=g).cuda(), torch.randint(0, 10, (8,), generator=g).cuda()
        la, lb = sa(x, y).item(), sb(x, y).item()
        assert abs(la - lb) <= 2e-3 * max(1.0, abs(lb)), (la, lb)
    for (n, pa), (_, pb) in zip(a.named_parameters(), b.named_parameters()):
        torch.testing.assert_close(pa, pb, rtol=2e-3, atol=2e-4, msg=lambda m, n=n: "%s: %s" % (n, m))


def test_resnet18_modules_train_step_matches_torchvision():
    import torchvision

    torch.manual_seed(0)
    ours = pdl.models.resnet18(num_classes=10).cuda()
    ref = torchvision.models.resnet18(num_classes=10).cuda()
    ref.load_state_dict(ours.state_dict())
    x, y = torch.randn(8, 3, 32, 32).cuda(), torch.randint(0, 10, (8,)).cuda()
    lo = pdl.nn.CrossEntropyLoss()(ours(x), y)
    lr = F.cross_entropy(ref(x), y)
    torch.testing.assert_close(lo, lr, rtol=1e-3, atol=1e-3)
    lo.backward()
    lr.backward()
    ga = dict(ours.named_parameters())["conv1.weight"].grad
    gb = dict(ref.named_parameters())["conv1.weight"].grad
    # batch 8 at 1x1 spatial extent makes the late BatchNorms ill-conditioned: compare directions, not ulps
    cos = torch.nn.functional.cosine_similarity(ga.flatten(), gb.flatten(), dim=0).item()
    assert cos > 0.99, cos
    for (n, a), (_, b) in zip(ours.named_buffers(), ref.named_buffers()):
        torch.testing.assert_close(a.float(), b.float(), rtol=1e-3, atol=1e-4, msg=lambda m, n=n: "%s: %s" % (n, m))


def test_ddp_unfused_optimizer_lets_the_user_edit_gradients():
    """``fused_optimizer=False``: the bucket kernel only averages, so what the reference's loop leaves room for between
    ``backward()`` and ``step()`` (launch_dist.py:91-92) -- here gradient clipping -- takes effect exactly as with torch."""
    dist = pdl.distributed
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % free_port(), world_size=1, rank=0)
    try:
        torch.manual_seed(0)
        model = pdl.models.ConvNet().cuda()
        ref = RefConvNet().cuda()
        ref.load_state_dict(copy.deepcopy(model.state_dict()))
        kw = dict(lr=0.05, momentum=0.9)
        opt, ropt = pdl.optim.SGD(model.parameters(), **kw), torch.optim.SGD(ref.parameters(), **kw)
        ddp = pdl.nn.parallel.DistributedDataParallel(model, device_ids=[0], fused_optimizer=False)
        crit = pdl.nn.CrossEntropyLoss()
        g = torch.Generator().manual_seed(7)
        for _ in range(3):
            x, y = torch.randn(12, 1, 28, 28, generator=g).cuda(), torch.randint(0, 10, (12,), generator=g).cuda()
            before = [p.detach().clone() for p in model.parameters()]
            loss = crit(ddp(x), y)
            opt.zero_grad()
            loss.backward()
            torch.cuda.synchronize()
            assert all(torch.equal(a, p.detach()) for a, p in zip(before, model.parameters())), "backward() must not update the weights"
            torch.nn.utils.clip_grad_norm_(model.parameters(), 0.05)
            opt.step()
            rl = F.cross_entropy(ref(x), y)
            ropt.zero_grad()
            rl.backward()
            torch.nn.utils.clip_grad_norm_(ref.parameters(), 0.05)
            ropt.step()
        assert ddp.get_step_stats()["fused_buckets"] == 0
        for a, b in zip(model.parameters(), ref.parameters()):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-5)
        step = pdl.TrainStep(ddp, crit, opt)  # still graphable: the captured iteration ends with the stand-alone SGD kernel
        assert step.graphed and type(step.engine).__name__ == "GenericGraphEngine"
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("wrapped", [False, True])
def test_resume_from_optimizer_state_dict_matches_uninterrupted_run(wrapped):
    """Momentum lives in a flat arena that the fused kernels update; ``load_state_dict`` must refill it: a run
    that is checkpointed after 3 steps and resumed in fresh objects ends where the uninterrupted run ends."""
    dist = pdl.distributed
    if wrapped:
        dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % free_port(), world_size=1, rank=0)
    try:
        kw = dict(lr=0.05, momentum=0.9, weight_decay=1e-4, nesterov=True)
        g = torch.Generator().manual_seed(11)
        batches = [(torch.randn(16, 1, 28, 28, generator=g).cuda(), torch.randint(0, 10, (16,), generator=g).cuda()) for _ in range(6)]

        def make(state=None):
            torch.manual_seed(0)
            m = pdl.models.ConvNet().cuda()
            if state is not None:
                m.load_state_dict(state)
            o = pdl.optim.SGD(m.parameters(), **kw)
            w = pdl.nn.parallel.DistributedDataParallel(m, device_ids=[0]) if wrapped else m
            return m, o, pdl.TrainStep(w, pdl.nn.CrossEntropyLoss(), o)

        m_full, _, s_full = make()
        for x, y in batches:
            s_full(x, y)
        m_a, o_a, s_a = make()
        for x, y in batches[:3]:
            s_a(x, y)
        torch.cuda.synchronize()
        ckpt_model, ckpt_opt = copy.deepcopy(m_a.state_dict()), copy.deepcopy(o_a.state_dict())
        ckpt_opt["param_groups"][0]["lr"] = kw["lr"]
        m_b, o_b, s_b = make(ckpt_model)
        s_b(*batches[0])  # builds the graph and the arenas BEFORE the load (the harder order); then restore both
        m_b.load_state_dict(ckpt_model)
        o_b.load_state_dict(ckpt_opt)
        for x, y in batches[3:]:
            s_b(x, y)
        torch.cuda.synchronize()
        for (n, a), (_, b) in zip(m_full.named_parameters(), m_b.named_parameters()):
            torch.testing.assert_close(a, b, rtol=1e-5, atol=1e-6, msg=lambda m, n=n: "%s: %s" % (n, m))
    finally:
        if wrapped:
            dist.destroy_process_group()


def test_six_call_loop_runs_on_the_graphed_step():
    """The reference's loop body, unchanged (launch_dist.py:86-92), on a wrapped ConvNet: forward is deferred, ``loss.backward()``
    replays the whole-iteration graph, ``optimizer.step()`` finds the update done.  Must follow the same trajectory as an explicit
    ``TrainStep`` and as torch; ``loss.item()`` and the outputs are readable afterwards; touching the outputs BEFORE backward
    falls back to the ordinary eager pass for that iteration."""
    from pytorch_distributed_learning_b200.parallel.lazy_step import DeferredLoss, DeferredOutput

    dist = pdl.distributed
    dist.init_process_group("nccl", init_method="tcp://127.0.0.1:%d" % free_port(), world_size=1, rank=0)
    try:
        torch.manual_seed(0)
        model = pdl.models.ConvNet().cuda()
        state = copy.deepcopy(model.state_dict())
        kw = dict(lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)
        opt = pdl.optim.SGD(model.parameters(), **kw)
        ddp = pdl.nn.parallel.DistributedDataParallel(model, device_ids=[0], output_device=0)
        crit = pdl.nn.CrossEntropyLoss().cuda(0)
        g = torch.Generator().manual_seed(1)
        batches = [(torch.randn(12, 1, 28, 28, generator=g), torch.randint(0, 10, (12,), generator=g)) for _ in range(5)]
        losses, kinds = [], []
        for i, (x, y) in enumerate(batches):
            images, labels = x.cuda(non_blocking=True), y.cuda(non_blocking=True)
            outputs = ddp(images)
            if i == 2:
                assert outputs.argmax(dim=1).shape == (12,)  # needs real logits now: this iteration runs eagerly
            loss = crit(outputs, labels)
            opt.zero_grad()
            loss.backward()
            opt.step()
            kinds.append((type(outputs).__name__, type(loss).__name__, getattr(loss, "_engine", None) is not None))
            losses.append(loss.item())
            assert torch.max(outputs, 1)[1].shape == (12,)  # after the step: the graph's logits (or the eager tensor)
            assert model.conv1.weight.grad is not None
        assert kinds[0] == ("DeferredOutput", "DeferredLoss", True) and kinds[1][2] and kinds[3][2] and kinds[4][2]
        assert kinds[2][2] is False, "an iteration whose outputs were used before backward() must not replay the graph"
        assert isinstance(outputs, DeferredOutput) and isinstance(loss, DeferredLoss)
        ref, ref_losses = _reference_run(state, batches, kw)
        torch.testing.assert_close(torch.tensor(losses), torch.tensor(ref_losses), rtol=1e-4, atol=1e-5)
        for (n, a), (_, b) in zip(model.named_parameters(), ref.named_parameters()):
            torch.testing.assert_close(a, b, rtol=1e-4, atol=5e-5, msg=lambda m, n=n: "%s: %s" % (n, m))
        # evaluation and no_grad calls are ordinary tensors
        ddp.eval()
        with torch.no_grad():
            assert isinstance(ddp(batches[0][0].cuda()), torch.Tensor)
    finally:
        dist.destroy_process_group()
