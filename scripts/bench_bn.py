"""Per-shape timing of the fused BatchNorm kernels on ResNet-50's activation shapes.
Usage: bench_bn.py [batch] [iters].  Reports ms and the HBM bandwidth implied by the algorithmic bytes
(forward: read x twice, write y [+ read residual]; backward: read x, dy twice [+ y], write dx [+ dres])."""
import sys, torch
sys.path.insert(0, ".")
import pytorch_distributed_learning_b200 as pdl

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
IT = int(sys.argv[2]) if len(sys.argv) > 2 else 10
dev = torch.device("cuda", 0)
CL = torch.channels_last
# (C, H, count plain bn+relu, count bn+residual+relu, count bn only (downsample))
SHAPES = [(64, 112, 1, 0, 0), (64, 56, 6, 0, 0), (256, 56, 0, 3, 1), (128, 56, 1, 0, 0), (128, 28, 7, 0, 0), (512, 28, 0, 4, 1),
          (256, 28, 1, 0, 0), (256, 14, 11, 0, 0), (1024, 14, 0, 6, 1), (512, 14, 1, 0, 0), (512, 7, 5, 0, 0), (2048, 7, 0, 3, 1)]


def timed(fn):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(IT):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / IT


tot_f = tot_b = 0.0
print("batch %d  %-22s %9s %9s   GB/s fwd  GB/s bwd" % (B, "shape / form", "fwd ms", "bwd ms"))
for C, H, n_relu, n_res, n_plain in SHAPES:
    for form, cnt in (("relu", n_relu), ("res+relu", n_res), ("plain", n_plain)):
        if not cnt:
            continue
        bn = pdl.nn.BatchNorm2d(C).to(dev)
        x = torch.randn(B, C, H, H, device=dev).contiguous(memory_format=CL).requires_grad_(True)
        r = torch.randn(B, C, H, H, device=dev).contiguous(memory_format=CL).requires_grad_(True) if form == "res+relu" else None
        relu = form != "plain"
        y = bn(x, relu=relu, residual=r)
        dy = torch.randn_like(y)
        t_f = timed(lambda: bn(x, relu=relu, residual=r))
        t_b = timed(lambda: torch.autograd.grad(y, [x] + ([r] if r is not None else []), dy, retain_graph=True))
        unit = x.numel() * 4.0
        bytes_f = unit * (3 + (1 if r is not None else 0))
        bytes_b = unit * (5 + (3 if r is not None else 0))  # x, dy twice (+ y twice), dx (+ dres)
        print("  C=%4d %3dx%-3d %-9s x%-2d %9.3f %9.3f   %7.0f   %7.0f" % (C, H, H, form, cnt, t_f, t_b, bytes_f / t_f / 1e6, bytes_b / t_b / 1e6))
        tot_f += cnt * t_f
        tot_b += cnt * t_b
print("ResNet-50 BN totals (ms): fwd %.3f  bwd %.3f" % (tot_f, tot_b))
