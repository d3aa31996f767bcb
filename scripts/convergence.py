"""Does it train?  A learnable synthetic task (ten noisy class prototypes), same initial weights and same
batches for this framework's graphed TrainStep and for the reference's stack (stock torch modules + torch.optim.SGD),
loss / accuracy curves side by side.  Usage: convergence.py [convnet|resnet18] [steps]"""
import copy
import sys

import torch
import torch.nn.functional as F

sys.path.insert(0, ".")
sys.path.insert(0, "tests")
import pytorch_distributed_learning_b200 as pdl

name = sys.argv[1] if len(sys.argv) > 1 else "convnet"
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 400
dev = torch.device("cuda", 0)
torch.manual_seed(0)
if name == "convnet":
    from test_gpu_engine import RefConvNet

    ours, ref, shape, batch = pdl.models.ConvNet().to(dev), RefConvNet().to(dev), (1, 28, 28), 100
    kw = dict(lr=0.05, momentum=0.9)
else:
    import torchvision

    ours, ref, shape, batch = pdl.models.resnet18(num_classes=10).to(dev), torchvision.models.resnet18(num_classes=10).to(dev), (3, 32, 32), 128
    kw = dict(lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)
ref.load_state_dict(copy.deepcopy(ours.state_dict()))
g = torch.Generator().manual_seed(1)
protos = torch.randn((10,) + shape, generator=g)  # ten class prototypes; a sample is its prototype plus unit noise
pool = 64
ys = [torch.randint(0, 10, (batch,), generator=g) for _ in range(pool)]
xs = [0.5 * protos[y] + torch.randn((batch,) + shape, generator=g) for y in ys]
xs = [x.to(dev) for x in xs]
ys = [y.to(dev) for y in ys]
oa, ob = pdl.optim.SGD(ours.parameters(), **kw), torch.optim.SGD(ref.parameters(), **kw)
step = pdl.TrainStep(ours, pdl.nn.CrossEntropyLoss(), oa)
print("%s, batch %d, %d steps, SGD %s; graphed=%s" % (name, batch, steps, kw, step.graphed))
print("step | ours loss | torch loss")
la = lb = None
for it in range(steps):
    x, y = xs[it % pool], ys[it % pool]
    la = step(x, y)
    out = ref(x)
    lb = F.cross_entropy(out, y)
    ob.zero_grad()
    lb.backward()
    ob.step()
    if it % (steps // 10) == 0 or it == steps - 1:
        print("%4d | %.4f | %.4f" % (it, la.item(), lb.item()))
ours.eval(), ref.eval()
with torch.no_grad():
    acc_a = sum((ours(x).argmax(1) == y).float().mean().item() for x, y in zip(xs, ys)) / pool
    acc_b = sum((ref(x).argmax(1) == y).float().mean().item() for x, y in zip(xs, ys)) / pool
print("train-set accuracy: ours %.3f, torch %.3f" % (acc_a, acc_b))
