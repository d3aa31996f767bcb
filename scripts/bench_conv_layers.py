"""Per-layer timing of the tcgen05 convolution kernels on ResNet-50's distinct layer shapes (forward, dgrad, wgrad).
Usage: bench_conv_layers.py [batch] [iters].  CUDA events around `iters` back-to-back launches after 3 warm-ups;
reports ms, TFLOP/s and the bandwidth implied by the algorithmic bytes (inputs + outputs once)."""
import sys, torch
sys.path.insert(0, ".")
from pytorch_distributed_learning_b200.ops import tc_conv

B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
IT = int(sys.argv[2]) if len(sys.argv) > 2 else 10
torch.backends.cudnn.allow_tf32 = True
dev = torch.device("cuda", 0)
CL = torch.channels_last
# (Cin, Cout, k, stride, H, count in ResNet-50)
LAYERS = [(64, 64, 1, 1, 56, 1), (64, 64, 3, 1, 56, 3), (64, 256, 1, 1, 56, 4), (256, 64, 1, 1, 56, 2), (256, 128, 1, 1, 56, 1),
          (128, 128, 3, 2, 56, 1), (128, 512, 1, 1, 28, 4), (256, 512, 1, 2, 56, 1), (512, 128, 1, 1, 28, 3), (128, 128, 3, 1, 28, 3),
          (512, 256, 1, 1, 28, 1), (256, 256, 3, 2, 28, 1), (256, 1024, 1, 1, 14, 6), (512, 1024, 1, 2, 28, 1), (1024, 256, 1, 1, 14, 5),
          (256, 256, 3, 1, 14, 5), (1024, 512, 1, 1, 14, 1), (512, 512, 3, 2, 14, 1), (512, 2048, 1, 1, 7, 3), (1024, 2048, 1, 2, 14, 1),
          (2048, 512, 1, 1, 7, 2), (512, 512, 3, 1, 7, 2)]


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


tot = {"fwd": 0.0, "dgrad": 0.0, "wgrad": 0.0}
print("batch %d   %-28s %8s %8s %8s   TFLOP/s fwd/dgrad/wgrad   GB/s(alg) fwd/dgrad/wgrad" % (B, "layer", "fwd ms", "dgrad ms", "wgrad ms"))
for cin, cout, k, s, h, cnt in LAYERS:
    pad = k // 2
    p = (h + 2 * pad - k) // s + 1
    x = torch.randn(B, cin, h, h, device=dev).contiguous(memory_format=CL)
    w = torch.randn(cout, cin, k, k, device=dev) * 0.05
    y = torch.empty(B, cout, p, p, device=dev).contiguous(memory_format=CL)
    dy = torch.randn_like(y)
    dx = torch.empty_like(x)
    dw = torch.empty_like(w)
    t_f = timed(lambda: tc_conv.conv2d_fwd(x, w, None, y, s, pad))
    t_d = timed(lambda: tc_conv.conv2d_dgrad(dy, w, dx, s, pad))
    t_w = timed(lambda: tc_conv.conv2d_wgrad(x, dy, dw, None, s, pad))
    tc_conv.check_errors(dev)
    flops = 2.0 * y.numel() * cin * k * k
    byts = (x.numel() + y.numel()) * 4.0
    print("  %4d->%4d k%d s%d @%3d x%d   %8.3f %8.3f %8.3f   %6.0f %6.0f %6.0f   %6.0f %6.0f %6.0f" % (
        cin, cout, k, s, h, cnt, t_f, t_d, t_w, flops / t_f / 1e9, flops / t_d / 1e9, flops / t_w / 1e9,
        byts / t_f / 1e6, byts / t_d / 1e6, byts / t_w / 1e6))
    tot["fwd"] += cnt * t_f; tot["dgrad"] += cnt * t_d; tot["wgrad"] += cnt * t_w
print("ResNet-50 conv totals (ms, weighted by layer count, stem excluded):", {k: round(v, 3) for k, v in tot.items()})
