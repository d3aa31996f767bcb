// Any-shape Conv2d forward / dgrad / wgrad on CUDA cores (strict fp32), NHWC activations and
// standard [Cout][Cin][kh][kw] weights.  Covers every convolution of the reference's models
// (ConvNet: launch_dist.py:15-25; ResNet convs incl. 7x7/2 and strided 1x1) for which the
// tcgen05 path has no specialisation, and is the oracle the tensor-core kernels are tested
// against.  wgrad is split-K over N*P*Q into a scratch buffer followed by a fixed-order
// second-pass reduction (deterministic), with the bias gradient folded in as one extra column.
#include <atomic>

#include "igemm_simt.cuh"

namespace {

struct ConvShape {
  int N, H, W, Cin, Cout, KH, KW, stride, pad, P, Q;
};

struct FwdA {
  const float* x; ConvShape s;
  __device__ float operator()(int m, int k) const {
    int q = m % s.Q, t = m / s.Q, p = t % s.P, n = t / s.P;
    int ci = k % s.Cin, rs = k / s.Cin, sx = rs % s.KW, r = rs / s.KW;
    int h = p * s.stride - s.pad + r, w = q * s.stride - s.pad + sx;
    if (h < 0 || h >= s.H || w < 0 || w >= s.W) return 0.0f;
    return x[((static_cast<long long>(n) * s.H + h) * s.W + w) * s.Cin + ci];
  }
};
struct FwdB {
  const float* w; ConvShape s;
  __device__ float operator()(int k, int co) const {
    int ci = k % s.Cin, rs = k / s.Cin;
    return w[(static_cast<long long>(co) * s.Cin + ci) * (s.KH * s.KW) + rs];
  }
};
struct FwdC {
  float* y; const float* bias; int Cout; int relu;
  __device__ void operator()(int m, int n, float v, int) const {
    if (bias) v += bias[n];
    if (relu) v = fmaxf(v, 0.0f);
    y[static_cast<long long>(m) * Cout + n] = v;
  }
};

struct DgradA {
  const float* dy; ConvShape s;
  __device__ float operator()(int m, int k) const {
    int w = m % s.W, t = m / s.W, h = t % s.H, n = t / s.H;
    int co = k % s.Cout, rs = k / s.Cout, sx = rs % s.KW, r = rs / s.KW;
    int ph = h + s.pad - r, qw = w + s.pad - sx;
    if (ph < 0 || qw < 0 || ph % s.stride || qw % s.stride) return 0.0f;
    int p = ph / s.stride, q = qw / s.stride;
    if (p >= s.P || q >= s.Q) return 0.0f;
    return dy[((static_cast<long long>(n) * s.P + p) * s.Q + q) * s.Cout + co];
  }
};
struct DgradB {
  const float* w; ConvShape s;
  __device__ float operator()(int k, int ci) const {
    int co = k % s.Cout, rs = k / s.Cout;
    return w[(static_cast<long long>(co) * s.Cin + ci) * (s.KH * s.KW) + rs];
  }
};
struct DgradC {
  float* dx; int Cin;
  __device__ void operator()(int m, int n, float v, int) const { dx[static_cast<long long>(m) * Cin + n] = v; }
};

// wgrad GEMM: M = Cout, N = Cin*KH*KW (+1 bias column), K = N*P*Q
struct WgradA {
  const float* dy; int Cout;
  __device__ float operator()(int co, int k) const { return dy[static_cast<long long>(k) * Cout + co]; }
};
struct WgradB {
  const float* x; ConvShape s; int ncols;  // column ncols-1 (when bias) is all ones
  bool bias;
  __device__ float operator()(int k, int j) const {
    if (bias && j == ncols - 1) return 1.0f;
    int rs = j % (s.KH * s.KW), ci = j / (s.KH * s.KW), sx = rs % s.KW, r = rs / s.KW;
    int q = k % s.Q, t = k / s.Q, p = t % s.P, n = t / s.P;
    int h = p * s.stride - s.pad + r, w = q * s.stride - s.pad + sx;
    if (h < 0 || h >= s.H || w < 0 || w >= s.W) return 0.0f;
    return x[((static_cast<long long>(n) * s.H + h) * s.W + w) * s.Cin + ci];
  }
};
struct WgradC {
  float* partial; int ncols; long long slab;  // partial[split][Cout][ncols]
  __device__ void operator()(int co, int j, float v, int split) const { partial[split * slab + static_cast<long long>(co) * ncols + j] = v; }
};

__global__ void wgrad_reduce_kernel(const float* __restrict__ partial, int splits, int Cout, int ncols, bool bias, float* __restrict__ dw,
                                    float* __restrict__ db) {
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  long long total = static_cast<long long>(Cout) * ncols;
  if (i >= total) return;
  float acc = 0.0f;
  for (int s = 0; s < splits; ++s) acc += partial[s * total + i];  // fixed order: deterministic
  int co = static_cast<int>(i / ncols), j = static_cast<int>(i % ncols);
  if (bias && j == ncols - 1) db[co] = acc;
  else dw[static_cast<long long>(co) * (ncols - (bias ? 1 : 0)) + j] = acc;
}

ConvShape make_shape(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad) {
  ConvShape s{N, H, W, Cin, Cout, KH, KW, stride, pad, 0, 0};
  s.P = (H + 2 * pad - KH) / stride + 1;
  s.Q = (W + 2 * pad - KW) / stride + 1;
  return s;
}

int wgrad_splits(const ConvShape& s) {
  long long K = static_cast<long long>(s.N) * s.P * s.Q;
  int tiles = pdl_ceil_div(s.Cout, simt::BM) * pdl_ceil_div(s.Cin * s.KH * s.KW + 1, simt::BN);
  int want = (148 * 4 + tiles - 1) / tiles;  // ~4 CTAs per SM
  long long max_by_k = (K + 255) / 256;
  if (want > max_by_k) want = static_cast<int>(max_by_k);
  if (want > 256) want = 256;
  if (want < 1) want = 1;
  return simt::num_splits(static_cast<int>(K), want);
}

}  // namespace

PDL_API int pdl_conv2d_fwd_simt(const float* x, const float* w, const float* bias, float* y, int N, int H, int W, int Cin, int Cout, int KH,
                                int KW, int stride, int pad, int relu, cudaStream_t stream) {
  ConvShape s = make_shape(N, H, W, Cin, Cout, KH, KW, stride, pad);
  return simt::launch(N * s.P * s.Q, Cout, KH * KW * Cin, 1, FwdA{x, s}, FwdB{w, s}, FwdC{y, bias, Cout, relu}, stream);
}

PDL_API int pdl_conv2d_dgrad_simt(const float* dy, const float* w, float* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                                  int stride, int pad, cudaStream_t stream) {
  ConvShape s = make_shape(N, H, W, Cin, Cout, KH, KW, stride, pad);
  return simt::launch(N * H * W, Cin, KH * KW * Cout, 1, DgradA{dy, s}, DgradB{w, s}, DgradC{dx, Cin}, stream);
}

// Bytes of scratch pdl_conv2d_wgrad_simt needs for this shape (split-K partial sums).
PDL_API long long pdl_conv2d_wgrad_scratch_bytes(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad) {
  ConvShape s = make_shape(N, H, W, Cin, Cout, KH, KW, stride, pad);
  return static_cast<long long>(wgrad_splits(s)) * Cout * (Cin * KH * KW + 1) * sizeof(float);
}

// dw[Cout][Cin][KH][KW] and (optional) db[Cout]; `scratch` must hold pdl_conv2d_wgrad_scratch_bytes().
PDL_API int pdl_conv2d_wgrad_simt(const float* x, const float* dy, float* dw, float* db, float* scratch, int N, int H, int W, int Cin,
                                  int Cout, int KH, int KW, int stride, int pad, cudaStream_t stream) {
  ConvShape s = make_shape(N, H, W, Cin, Cout, KH, KW, stride, pad);
  const bool bias = db != nullptr;
  const int ncols = Cin * KH * KW + (bias ? 1 : 0);
  const int K = N * s.P * s.Q;
  const int splits = wgrad_splits(s);
  const long long slab = static_cast<long long>(Cout) * ncols;
  if (int rc = simt::launch(Cout, ncols, K, splits, WgradA{dy, Cout}, WgradB{x, s, ncols, bias}, WgradC{scratch, ncols, slab}, stream)) return rc;
  int threads = 256;
  wgrad_reduce_kernel<<<pdl_ceil_div(slab, threads), threads, 0, stream>>>(scratch, splits, Cout, ncols, bias, dw, db);
  PDL_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------------
// Weight gradient for callers that are not allowed to allocate.
//
// Some hosts of this library cannot obtain device memory at the point where they need a wgrad:
// processes in which another allocator (a framework's caching pool, a serving runtime's arena) has
// already reserved the whole device, so a fresh cudaMalloc simply fails; and latency-critical loops
// that must not call an allocator at all because cudaMalloc/cudaFree may synchronise the device.
// For them the split-K partial sums live in a fixed array inside the library image instead of a
// caller-provided buffer.  The array is kNoAllocScratchBytes (8 MiB) and every GPU that loads the
// library carries it until the process exits; shapes whose partial sums do not fit are refused.
//
// The array is shared by every call in the process: two calls whose kernels can overlap on the GPU
// (different streams) would overwrite each other's partial sums, so callers must keep such calls
// ordered.  Nothing needs clearing between calls: every partial-sum slot that the reduction pass
// reads is written by the GEMM pass of the same call first (leftovers of the previous call are
// never read).  The first call on a device looks up the device address of the array
// (cudaGetSymbolAddress); make that first call before any stream capture begins.
// ------------------------------------------------------------------------------------------------
namespace {
constexpr long long kNoAllocScratchBytes = 8ll << 20;
constexpr int kNoAllocMaxDevices = 64;
}  // namespace

static __device__ __align__(256) float g_wgrad_noalloc_scratch[kNoAllocScratchBytes / sizeof(float)];

namespace {
// Device address of g_wgrad_noalloc_scratch per device ordinal.  In host code the symbol's own name is only
// a host-side shadow, not a device pointer, so it must never be handed to a kernel directly.
// Atomic: two host threads may make the first call on a device at the same time; both resolve the same address.
std::atomic<float*> g_noalloc_dev_ptr[kNoAllocMaxDevices];
}  // namespace

// Largest scratch requirement (bytes) pdl_conv2d_wgrad_simt_noalloc accepts.
PDL_API long long pdl_conv2d_wgrad_noalloc_max_bytes() { return kNoAllocScratchBytes; }

// Same result as pdl_conv2d_wgrad_simt, but takes no scratch pointer and performs no allocation.
PDL_API int pdl_conv2d_wgrad_simt_noalloc(const float* x, const float* dy, float* dw, float* db, int N, int H, int W, int Cin, int Cout, int KH,
                                          int KW, int stride, int pad, cudaStream_t stream) {
  const long long need = pdl_conv2d_wgrad_scratch_bytes(N, H, W, Cin, Cout, KH, KW, stride, pad);
  if (need > kNoAllocScratchBytes) {
    pdl_set_error("wgrad (no-alloc): this shape needs %lld bytes of partial sums, the built-in buffer holds %lld; "
                  "use pdl_conv2d_wgrad_simt with caller-provided scratch", need, kNoAllocScratchBytes);
    return -10;
  }
  int dev = -1;
  PDL_CUDA_CHECK(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kNoAllocMaxDevices) { pdl_set_error("wgrad (no-alloc): device ordinal %d out of range", dev); return -11; }
  float* scratch = g_noalloc_dev_ptr[dev].load(std::memory_order_acquire);
  if (scratch == nullptr) {  // once per device, with that device current
    PDL_CUDA_CHECK(cudaGetSymbolAddress(reinterpret_cast<void**>(&scratch), g_wgrad_noalloc_scratch));
    g_noalloc_dev_ptr[dev].store(scratch, std::memory_order_release);
  }
  return pdl_conv2d_wgrad_simt(x, dy, dw, db, scratch, N, H, W, Cin, Cout, KH, KW, stride, pad, stream);
}
