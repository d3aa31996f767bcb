"""The C++ reducer state machine (csrc/host/reducer.cpp): in-order bucket launches, tail flush, arrival order, misuse."""
import pytest

from pytorch_distributed_learning_b200.parallel.reducer import ReducerCore


def test_buckets_launch_in_order_only_when_complete():
    # 6 parameters, 3 buckets: {5, 4} -> 0, {3, 2} -> 1, {1, 0} -> 2 (reverse registration order, as DDP assumes)
    core = ReducerCore(6, [2, 2, 1, 1, 0, 0])
    core.prepare(record_arrival=True)
    assert list(core.mark_ready(3)) == [] and list(core.mark_ready(2)) == []  # bucket 1 is complete but bucket 0 is not
    assert core.next_bucket == 0
    assert list(core.mark_ready(5)) == []
    assert list(core.mark_ready(4)) == [0, 1]  # bucket 0 completes and releases the waiting bucket 1 with it
    assert core.is_ready(4) and not core.is_ready(0)
    assert list(core.mark_ready(0)) == [] and list(core.mark_ready(1)) == [2]
    assert list(core.finalize()) == [] and core.arrival() == [3, 2, 5, 4, 0, 1]


def test_tail_flush_and_second_iteration():
    core = ReducerCore(4, [0, 0, 1, 1])
    core.prepare(record_arrival=True)
    assert list(core.mark_ready(1)) == []
    assert list(core.finalize()) == [0, 1]  # parameters 0, 2, 3 produced no gradient: their buckets are flushed in order
    first = core.arrival()
    core.prepare(record_arrival=False)  # later iterations do not disturb the recorded order
    assert list(core.mark_ready(0)) == [] and list(core.mark_ready(1)) == [0]
    assert list(core.mark_ready(3)) == [] and list(core.mark_ready(2)) == [1]
    assert core.arrival() == first == [1]


def test_misuse_is_reported():
    core = ReducerCore(3, [0, 0, 1])
    assert list(core.mark_ready(0)) == []  # no forward through the wrapper yet: not synchronised, nothing to launch
    assert list(core.finalize()) == []
    core.prepare()
    core.mark_ready(0)
    with pytest.raises(RuntimeError):
        core.mark_ready(0)  # second gradient for a parameter within one iteration
    with pytest.raises(IndexError):
        core.mark_ready(7)
    with pytest.raises(ValueError):
        ReducerCore(2, [0, -1])
    core.set_layout(3, [1, 0, 0])  # rebuild: a new layout starts clean
    core.prepare()
    assert list(core.mark_ready(0)) == [] and list(core.mark_ready(2)) == [] and list(core.mark_ready(1)) == [0, 1]
    core.close()


def test_ddp_rebuilds_buckets_from_the_observed_arrival_order():
    """A model that uses its layers in the opposite order of their registration, with a tiny bucket cap: after the first
    iteration DDP re-lays its buckets out in arrival order (once), and training still equals plain SGD (world size 1)."""
    import copy

    import torch

    import pytorch_distributed_learning_b200 as pdl
    from conftest import free_port

    class Backwards(torch.nn.Module):
        def __init__(self):
            super().__init__()
            self.a, self.b, self.c = torch.nn.Linear(8, 8), torch.nn.Linear(8, 8), torch.nn.Linear(8, 8)
            self.unused = torch.nn.Linear(8, 8)  # never takes part: its bucket is flushed with zeros at the end

        def forward(self, x):
            return self.a(torch.relu(self.b(torch.relu(self.c(x)))))  # registration order a, b, c; use order c, b, a

    dist = pdl.distributed
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % free_port(), world_size=1, rank=0)
    try:
        torch.manual_seed(0)
        ref = Backwards()
        model = copy.deepcopy(ref)
        ddp = pdl.DistributedDataParallel(model, bucket_cap_mb=200 / (1 << 20), first_bucket_mb=200 / (1 << 20))
        opt_a = torch.optim.SGD(ref.parameters(), lr=0.1, momentum=0.9)
        opt_b = pdl.optim.SGD(ddp.parameters(), lr=0.1, momentum=0.9)
        before = [list(bk.param_ids) for bk in ddp._buckets]
        for step in range(4):
            x = torch.randn(5, 8)
            for net, opt in ((ref, opt_a), (ddp, opt_b)):
                opt.zero_grad()
                net(x).square().mean().backward()
                opt.step()
        after = [list(bk.param_ids) for bk in ddp._buckets]
        assert ddp._stats["rebuilds"] == 1 and before != after
        # gradients arrive a, b, c (reverse of use); the never-used layer goes last
        names = [ddp._param_names[i] for bk in ddp._buckets for i in bk.param_ids]
        assert names[:2] == ["a.bias", "a.weight"] or names[:2] == ["a.weight", "a.bias"]
        assert set(names[-2:]) == {"unused.weight", "unused.bias"}
        for (n, p), (_, q) in zip(ref.named_parameters(), model.named_parameters()):
            if not n.startswith("unused"):
                torch.testing.assert_close(p, q, rtol=1e-6, atol=1e-7, msg=lambda m, n=n: "%s: %s" % (n, m))
    finally:
        dist.destroy_process_group()
