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
