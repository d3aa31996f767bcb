"""A couple of eager ResNet training steps on one GPU (target of the ncu launch list)."""
import sys, torch
sys.path.insert(0, ".")
import pytorch_distributed_learning_b200 as pdl
name = sys.argv[1] if len(sys.argv) > 1 else "resnet50"
N = int(sys.argv[2]) if len(sys.argv) > 2 else 32
res = 224 if name == "resnet50" else 32
torch.manual_seed(0)
model = getattr(pdl.models, name)(num_classes=1000 if name == "resnet50" else 10).cuda()
opt = pdl.optim.SGD(model.parameters(), lr=0.02, momentum=0.9, weight_decay=1e-4, nesterov=True)
crit = pdl.nn.CrossEntropyLoss()
x = torch.randn(N, 3, res, res).cuda(); y = torch.randint(0, 10, (N,)).cuda()
for _ in range(int(sys.argv[3]) if len(sys.argv) > 3 else 2):
    loss = crit(model(x), y); opt.zero_grad(); loss.backward(); opt.step()
torch.cuda.synchronize()
print("loss", loss.item())
