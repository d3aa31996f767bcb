"""Per-kernel timeline of ONE graphed ConvNet training step under DDP (torch.profiler / CUPTI on graph replays; works under
torchrun on several GPUs -- unlike ncu, which must never wrap a multi-rank command).  Shows which kernels overlap and what
sits on the critical path, e.g. whether the early gradient exchange on the branch stream is hidden behind the main chain.

    python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 scripts/prof_timeline.py [batch] [out.txt]
"""
import os
import sys

import torch

sys.path.insert(0, ".")
import pytorch_distributed_learning_b200 as pdl  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 100
out_path = sys.argv[2] if len(sys.argv) > 2 else None
rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
local = int(os.environ.get("LOCAL_RANK", str(rank)))
os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
os.environ.setdefault("MASTER_PORT", "29534")
os.environ.setdefault("RANK", "0")
os.environ.setdefault("WORLD_SIZE", "1")
torch.cuda.set_device(local)
dist = pdl.distributed
dist.init_process_group(backend="nccl")
torch.manual_seed(0)
model = pdl.models.ConvNet().cuda()
opt = pdl.optim.SGD(model.parameters(), lr=1e-4)
ddp = pdl.nn.parallel.DistributedDataParallel(model, device_ids=[local])
step = pdl.TrainStep(ddp, pdl.nn.CrossEntropyLoss(), opt)
x = torch.randn(N, 1, 28, 28).cuda()
y = torch.randint(0, 10, (N,)).cuda()
for _ in range(10):
    step(x, y)
torch.cuda.synchronize()
dist.barrier()
from torch.profiler import ProfilerActivity, profile  # noqa: E402

STEPS = 6
with profile(activities=[ProfilerActivity.CUDA]) as prof:
    dist._default().device_barrier()
    for _ in range(STEPS):
        step(x, y)
    torch.cuda.synchronize()
evs = []
for ev in prof.events():
    if ev.device_type is not None and "cuda" in str(ev.device_type).lower():
        nm = ev.name.replace("void ", "").replace("(anonymous namespace)::", "").replace("<unnamed>::", "").split("(")[0][:60]
        evs.append((ev.time_range.start, ev.time_range.end, nm))
evs.sort()
# split into steps at the input-refresh kernel (first kernel of every step)
starts = [i for i, e in enumerate(evs) if "copy_pair" in e[2]]
lines = []
if len(starts) >= STEPS:
    lo, hi = starts[-2], starts[-1]  # the second-to-last step: complete, warm
    t0 = evs[lo][0]
    lines.append("rank %d/%d batch %d: one step = %.1f us (from its first kernel to the next step's first kernel)" % (rank, world, N, evs[hi][0] - t0))
    lines.append("%9s %9s %8s  %s" % ("start_us", "end_us", "dur_us", "kernel"))
    for s, e, nm in evs[lo:hi]:
        lines.append("%9.1f %9.1f %8.1f  %s" % (s - t0, e - t0, e - s, nm))
text = "\n".join(lines)
everyone = dist._default().allgather_object(text)
if rank == 0:
    full = "\n\n".join(everyone)
    print(full)
    if out_path:
        with open(out_path, "w") as f:
            f.write(full + "\n")
dist.destroy_process_group()
