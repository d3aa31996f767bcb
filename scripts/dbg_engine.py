import sys, copy, torch
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import torch.nn.functional as F
import pytorch_distributed_learning_b200 as pdl
from test_gpu_engine import RefConvNet
torch.manual_seed(0)
N = int(sys.argv[1]) if len(sys.argv) > 1 else 100
model = pdl.models.ConvNet().cuda()
state = copy.deepcopy(model.state_dict())
opt = pdl.optim.SGD(model.parameters(), lr=1.0)
step = pdl.TrainStep(model, pdl.nn.CrossEntropyLoss(), opt)
x = torch.randn(N, 1, 28, 28).cuda(); y = torch.randint(0, 10, (N,)).cuda()
loss = step(x, y).item()
print("use_tc", step.engine.use_tc, "loss", loss, "kernels", step.kernels_per_step())
torch.backends.cudnn.allow_tf32 = False
ref = RefConvNet().cuda(); ref.load_state_dict(state)
l = F.cross_entropy(ref(x), y); l.backward()
print("ref loss", l.item())
for (n, p), (_, q) in zip(model.named_parameters(), ref.named_parameters()):
    g_ours = state[n].cuda() - p.detach()
    g = q.grad
    print("%-14s |g| %.3e  rel err %.3e  cos %.6f" % (n, g.norm().item(), ((g_ours - g).norm() / g.norm()).item(),
          F.cosine_similarity(g_ours.flatten(), g.flatten(), dim=0).item()))

torch.backends.cudnn.allow_tf32 = True
ref2 = RefConvNet().cuda(); ref2.load_state_dict(state)
F.cross_entropy(ref2(x), y).backward()
print("--- torch's own TF32 convs (cuDNN) vs its strict fp32:")
for (n, p), (_, q) in zip(ref2.named_parameters(), ref.named_parameters()):
    print("%-14s rel err %.3e" % (n, ((p.grad - q.grad).norm() / q.grad.norm()).item()))
