"""Replay the graphed ConvNet step many times from the same state (lr = 0) and report every static buffer
whose contents are not bitwise identical across replays: the first one in dataflow order names the flaky kernel.
Usage: dbg_determinism.py [batch] [strict|tc] [replays]"""
import sys, torch
sys.path.insert(0, ".")
import pytorch_distributed_learning_b200 as pdl

N = int(sys.argv[1]) if len(sys.argv) > 1 else 100
strict = (sys.argv[2] if len(sys.argv) > 2 else "strict") == "strict"
reps = int(sys.argv[3]) if len(sys.argv) > 3 else 200
torch.backends.cudnn.allow_tf32 = not strict
torch.backends.cuda.matmul.allow_tf32 = not strict
torch.manual_seed(0)
model = pdl.models.ConvNet().cuda()
opt = pdl.optim.SGD(model.parameters(), lr=0.0)
step = pdl.TrainStep(model, pdl.nn.CrossEntropyLoss(), opt)
x = torch.randn(N, 1, 28, 28).pin_memory(); y = torch.randint(0, 10, (N,)).pin_memory()
step(x, y); torch.cuda.synchronize()
eng = step.engine
entry = eng.graphs[eng.last_key]
bufs = {k: v for k, v in entry["buf"].items() if torch.is_tensor(v)}
for i, g in enumerate(eng.g_views):
    bufs["grad%d" % i] = g
ref = {k: v.clone() for k, v in bufs.items()}
bad = {}
for r in range(reps):
    step(x, y); torch.cuda.synchronize()
    for k, v in bufs.items():
        if k in ("scratch",):
            continue
        if not torch.equal(v, ref[k]):
            d = (v.float() - ref[k].float()).abs().max().item()
            bad.setdefault(k, [0, 0.0])
            bad[k][0] += 1; bad[k][1] = max(bad[k][1], d)
print("batch", N, "strict" if strict else "tc", "use_tc", eng.use_tc, "replays", reps)
print("buffers:", sorted(bufs))
print("non-deterministic:", {k: (c, "%.3e" % d) for k, (c, d) in bad.items()} or "none")
