import pytest
import torch

import pytorch_distributed_learning_b200 as pdl


@pytest.mark.parametrize("kw", [dict(), dict(momentum=0.9), dict(momentum=0.9, weight_decay=1e-4, nesterov=True),
                                dict(momentum=0.5, dampening=0.1, weight_decay=0.01)])
def test_cpu_matches_torch_sgd(kw):
    torch.manual_seed(0)
    a = [torch.randn(5, 3, requires_grad=True), torch.randn(7, requires_grad=True)]
    b = [t.detach().clone().requires_grad_(True) for t in a]
    oa, ob = pdl.optim.SGD(a, lr=0.02, **kw), torch.optim.SGD(b, lr=0.02, **kw)
    for step in range(4):
        for ta, tb in zip(a, b):
            g = torch.randn_like(ta)
            ta.grad, tb.grad = g.clone(), g.clone()
        oa.step()
        ob.step()
        for ta, tb in zip(a, b):
            torch.testing.assert_close(ta, tb, rtol=0, atol=0)
    sa, sb = oa.state_dict(), ob.state_dict()
    assert sa["param_groups"][0]["lr"] == sb["param_groups"][0]["lr"]
    assert set(sa["state"].keys()) == set(sb["state"].keys())


def test_registry_lookup_and_state_dict_roundtrip():
    w = torch.nn.Parameter(torch.zeros(3))
    other = torch.nn.Parameter(torch.zeros(3))
    opt = pdl.optim.SGD([w], lr=0.1, momentum=0.9)
    assert pdl.optim.lookup_optimizer(w)[0] is opt
    assert pdl.optim.lookup_optimizer(other) is None
    w.grad = torch.ones(3)
    opt.step()
    opt2 = pdl.optim.SGD([w], lr=0.5, momentum=0.9)
    opt2.load_state_dict(opt.state_dict())
    assert opt2.param_groups[0]["lr"] == 0.1
    assert torch.equal(opt2.state[w]["momentum_buffer"], torch.ones(3))


def test_load_state_dict_refills_flat_momentum_views_and_notifies_mirrors():
    """What the fused GPU path relies on: after load_state_dict the arena views are still the momentum buffers,
    hold the loaded values, and whoever mirrors hyper-parameters on the device is told to upload them again."""
    torch.manual_seed(0)
    shapes = [(5, 3), (7,)]

    def make():
        ps = [torch.nn.Parameter(torch.randn(*s)) for s in shapes]
        opt = pdl.optim.SGD(ps, lr=0.02, momentum=0.9, nesterov=True)
        flat = torch.zeros(sum(p.numel() for p in ps))
        off = 0
        for p in ps:
            opt.bind_momentum_view(p, flat[off:off + p.numel()].view_as(p))
            off += p.numel()
        return ps, opt, flat

    class Mirror:
        calls = []

        def on_reload(self, first):
            self.calls.append(first)

    pa, oa, flat_a = make()
    for _ in range(3):
        for p in pa:
            p.grad = torch.randn_like(p)
        oa.step()
    assert flat_a.abs().sum() > 0  # step() updated the arena through the views
    pb, ob, flat_b = make()
    mirror = Mirror()
    ob.add_reload_listener(mirror.on_reload)
    sd = oa.state_dict()
    sd["param_groups"][0]["lr"] = 0.5
    ob.load_state_dict(sd)
    assert torch.equal(flat_b, flat_a)
    for p in pb:
        assert ob.state[p]["momentum_buffer"].data_ptr() >= flat_b.data_ptr()
        assert ob.state[p]["momentum_buffer"].data_ptr() < flat_b.data_ptr() + flat_b.numel() * 4
    assert mirror.calls == [0]  # history present -> first_step = 0
    assert ob.group_by_id(pdl.optim.lookup_optimizer(pb[0])[1])["lr"] == 0.5  # groups are found by position after the reload
    # a state dict without history resets the views and says "first step"
    fresh = pdl.optim.SGD([torch.nn.Parameter(torch.randn(*s)) for s in shapes], lr=0.1, momentum=0.9).state_dict()
    ob.load_state_dict(fresh)
    assert mirror.calls == [0, 1] and flat_b.abs().sum() == 0


def test_deepcopy_and_pickle_give_a_working_optimizer():
    import copy
    import pickle

    w = torch.nn.Parameter(torch.ones(3))
    opt = pdl.optim.SGD([w], lr=0.1, momentum=0.9)
    w.grad = torch.ones(3)
    opt.step()
    for clone in (copy.deepcopy(opt), pickle.loads(pickle.dumps(opt))):
        p = clone.param_groups[0]["params"][0]
        assert p is not w and pdl.optim.lookup_optimizer(p)[0] is clone
        p.grad = torch.ones(3)
        clone.step()
        torch.testing.assert_close(p.detach(), torch.full((3,), 1 - 0.1 - 0.1 * 1.9))
