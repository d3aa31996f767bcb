import sys, torch
sys.path.insert(0, ".")
from pytorch_distributed_learning_b200.ops import tc_conv, functional as NF
from pytorch_distributed_learning_b200 import _lib
DEV = torch.device("cuda", 0)
N, Cin, H, Cout, k, pad = 5, 64, 10, 128, 3, 0
P = H - k + 1
torch.manual_seed(0)
x = torch.randn(N, Cin, H, H, device=DEV).contiguous(memory_format=torch.channels_last)
dy = torch.randn(N, Cout, P, P, device=DEV).contiguous(memory_format=torch.channels_last)
dw = torch.full((Cout, Cin, k, k), 7.0, device=DEV); db = torch.empty(Cout, device=DEV)
sc = NF.scratch(1 << 24, DEV, "tc_wgrad"); sc.fill_(0x7f)
print("ok", tc_conv.conv2d_wgrad(x, dy, dw, db, 1, pad))
torch.cuda.synchronize()
print("err flag", tc_conv.error_flag(DEV).item())
L = _lib.cuda()
nb = L.pdl_tc_wgrad_scratch_bytes(N, H, H, Cin, Cout, k, k, pad)
part = sc[:nb].view(torch.float32)
print("scratch bytes", nb, "partial absmax", part.abs().max().item(), "nonzero frac", (part != 0).float().mean().item())
print("dw absmax", dw.abs().max().item())
ref = torch.nn.grad.conv2d_weight(x, (Cout, Cin, k, k), dy)
print("ref absmax", ref.abs().max().item(), "err", (dw - ref).abs().max().item())
p = part.view(-1, 9, Cout, Cin)
print("splits", p.shape[0], "per-split absmax", [round(p[i].abs().max().item(), 2) for i in range(p.shape[0])])
tot = p.sum(0)  # [tap][co][ci]
refp = ref.permute(2, 3, 0, 1).reshape(9, Cout, Cin)
print("sum-of-partials vs ref err", (tot - refp).abs().max().item())
