"""Strict-fp32 graphed ConvNet step vs stock torch, per-parameter error after each of a few steps.
Usage: dbg_strict.py 100 7   (several batch sizes back to back in one process, like the test module)"""
import sys, copy, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import torch.nn.functional as F
import pytorch_distributed_learning_b200 as pdl
from test_gpu_engine import RefConvNet
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False


def run(N):
    torch.manual_seed(0)
    model = pdl.models.ConvNet().cuda()
    ref = RefConvNet().cuda(); ref.load_state_dict(copy.deepcopy(model.state_dict()))
    opt = pdl.optim.SGD(model.parameters(), lr=1e-2)
    ropt = torch.optim.SGD(ref.parameters(), lr=1e-2)
    step = pdl.TrainStep(model, pdl.nn.CrossEntropyLoss(), opt)
    g = torch.Generator().manual_seed(1)
    for it in range(5):
        x = torch.randn(N, 1, 28, 28, generator=g); y = torch.randint(0, 10, (N,), generator=g)
        before = {n: p.detach().clone() for n, p in model.named_parameters()}
        loss = step(x.pin_memory(), y.pin_memory()).item()
        ref64 = copy.deepcopy(ref).double()
        F.cross_entropy(ref64(x.cuda().double()), y.cuda()).backward()
        g64 = {n: q.grad for n, q in ref64.named_parameters()}
        l = F.cross_entropy(ref(x.cuda()), y.cuda()); ropt.zero_grad(); l.backward(); ropt.step()
        print("N", N, "it", it, "use_tc", step.engine.use_tc, "loss", loss, "ref", l.item())
        for (n, p), (_, q) in zip(model.named_parameters(), ref.named_parameters()):
            upd = (before[n] - p.detach()) / 1e-2
            err = (upd - q.grad).abs().max().item()
            if err > 1e-5:
                print("   %-14s |g|max %.3e  ours-torch %.3e   ours-fp64 %.3e   torch-fp64 %.3e" % (n, q.grad.abs().max().item(), err,
                      (upd.double() - g64[n]).abs().max().item(), (q.grad.double() - g64[n]).abs().max().item()))


for n in [int(a) for a in sys.argv[1:]] or [100, 7]:
    run(n)
