"""In-situ kernel timeline of the graphed training step (torch.profiler / CUPTI on graph replays): per-kernel totals
inside the CUDA graph, warm caches, real overlap -- unlike ncu's serialised cold-cache launch list.
Usage: prof_step.py [convnet|resnet18|resnet50] [batch] [steps]      (PDL_PROF_DDP=1: wrap the model in DistributedDataParallel,
world size 1, so that the reducer path -- gradient sinks, bucket kernels, fused SGD -- is part of the captured iteration)"""
import os
import collections
import sys

import torch

sys.path.insert(0, ".")
import pytorch_distributed_learning_b200 as pdl

name = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 64
steps = int(sys.argv[3]) if len(sys.argv) > 3 else 3
shape = {"convnet": (1, 28, 28), "resnet18": (3, 32, 32), "resnet50": (3, 224, 224)}[name]
classes = {"convnet": 10, "resnet18": 10, "resnet50": 1000}[name]
torch.manual_seed(0)
model = (pdl.models.ConvNet() if name == "convnet" else getattr(pdl.models, name)(num_classes=classes)).cuda()
kw = dict(lr=1e-4) if name == "convnet" else dict(lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)
opt = pdl.optim.SGD(model.parameters(), **kw)
if os.environ.get("PDL_PROF_DDP", "0") == "1":
    os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
    os.environ.setdefault("MASTER_PORT", "29541")
    pdl.distributed.init_process_group("nccl", world_size=1, rank=0)
    model = pdl.nn.parallel.DistributedDataParallel(model, device_ids=[0])
step = pdl.TrainStep(model, pdl.nn.CrossEntropyLoss(), opt)
x = torch.randn((N,) + shape).cuda()
y = torch.randint(0, classes, (N,)).cuda()
for _ in range(5):
    step(x, y)
torch.cuda.synchronize()
from torch.profiler import ProfilerActivity, profile

with profile(activities=[ProfilerActivity.CUDA]) as prof:
    for _ in range(steps):
        step(x, y)
    torch.cuda.synchronize()
agg = collections.defaultdict(lambda: [0, 0.0])
t_min, t_max = None, None
for ev in prof.events():
    if ev.device_type is not None and "cuda" in str(ev.device_type).lower():
        dur = ev.device_time_total if hasattr(ev, "device_time_total") else ev.cuda_time_total
        nm = ev.name.replace("void ", "").replace("(anonymous namespace)::", "").replace("<unnamed>::", "").split("(")[0][:70]
        agg[nm][0] += 1
        agg[nm][1] += dur
        t0 = ev.time_range.start
        t1 = ev.time_range.end
        t_min = t0 if t_min is None else min(t_min, t0)
        t_max = t1 if t_max is None else max(t_max, t1)
tot = sum(v[1] for v in agg.values())
print("%s batch %d: %d steps, kernel-time sum %.1f us/step, span %.1f us/step" % (name, N, steps, tot / steps, (t_max - t_min) / steps if t_min is not None else -1))
for nm, (c, t) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:28]:
    print("%9.1f us/step %5.1f%% %6.1f launches/step  avg %7.1f us  %s" % (t / steps, 100 * t / tot, c / steps, t / c, nm))
