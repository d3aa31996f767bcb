"""A few graphed ConvNet steps on one GPU (target of the ncu launch list / full capture)."""
import sys, torch
sys.path.insert(0, ".")
import pytorch_distributed_learning_b200 as pdl
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 3
torch.manual_seed(0)
model = pdl.models.ConvNet().cuda()
opt = pdl.optim.SGD(model.parameters(), lr=1e-4)
step = pdl.TrainStep(model, pdl.nn.CrossEntropyLoss(), opt)
x = torch.randn(N, 1, 28, 28).cuda(); y = torch.randint(0, 10, (N,)).cuda()
for _ in range(steps):
    loss = step(x, y)
torch.cuda.synchronize()
print("loss", loss.item(), "kernels/step", step.kernels_per_step())
