import sys, ctypes, torch
sys.path.insert(0, ".")
from pytorch_distributed_learning_b200.ops import tc_conv
from pytorch_distributed_learning_b200 import _lib
L = _lib.cuda(); L.pdl_tc_set_trace.argtypes = [ctypes.c_void_p]; L.pdl_tc_set_trace.restype = None
DEV = torch.device("cuda", 0)
tr = torch.zeros(16, dtype=torch.int64, device=DEV)
names = ["start", "setup done", "TMA ks0 issued", "TMA last issued", "MMA saw ks0", "MMA saw last", "epi saw acc", "epi done", "end", "first tmem_ld", "staging done", "after bar", "pool loop done"]
for (N, Cin, H, Cout, ps) in [(100, 32, 13, 64, 1), (100, 64, 10, 128, 2)]:
    x = torch.randn(N, Cin, H, H, device=DEV).contiguous(memory_format=torch.channels_last)
    w = torch.randn(Cout, Cin, 3, 3, device=DEV) * 0.1; b = torch.randn(Cout, device=DEV)
    for it in range(3):
        tc_conv.conv2d_fwd_relu_pool(x, w, b, ps)
    torch.cuda.synchronize()
    L.pdl_tc_set_trace(ctypes.c_void_p(tr.data_ptr()))
    tc_conv.conv2d_fwd_relu_pool(x, w, b, ps)
    torch.cuda.synchronize()
    L.pdl_tc_set_trace(None)
    t = tr.cpu().tolist()
    print("conv %d->%d pool/%d:" % (Cin, Cout, ps), ", ".join("%s +%d" % (n, t[i] - t[0]) for i, n in enumerate(names)))
