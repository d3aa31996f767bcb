// Element-wise / pooling / linear / loss kernels (NHWC fp32) for the reference's models:
//   ReLU, MaxPool2d incl. the overlapping 2/1 window (launch_dist.py:17-27), Linear 2048->10 in
//   strict fp32 (launch_dist.py:32; the reference's cuBLAS sgemm is not TF32), CrossEntropyLoss
//   mean-reduced with integer targets (launch_dist.py:58), ResNet's residual add+ReLU and
//   global average pool.  Pool backward is gather-style (each input looks at the <=ceil(k/s)^2
//   windows that contain it), so overlapping windows need no atomics and stay deterministic.
#include "igemm_simt.cuh"

#include <cfloat>

namespace {

constexpr int kThreads = 256;

__global__ void relu_fwd_kernel(const float4* __restrict__ x, float4* __restrict__ y, long long n4) {
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i < n4) {
    float4 v = x[i];
    y[i] = make_float4(fmaxf(v.x, 0.f), fmaxf(v.y, 0.f), fmaxf(v.z, 0.f), fmaxf(v.w, 0.f));
  }
}
__global__ void relu_fwd_tail(const float* __restrict__ x, float* __restrict__ y, long long from, long long n) {
  long long i = from + blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i < n) y[i] = fmaxf(x[i], 0.f);
}
__global__ void relu_bwd_kernel(const float* __restrict__ y, const float* __restrict__ dy, float* __restrict__ dx, long long n) {
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i < n) dx[i] = y[i] > 0.f ? dy[i] : 0.f;
}
__global__ void add_relu_fwd_kernel(const float* __restrict__ a, const float* __restrict__ b, float* __restrict__ y, long long n) {
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i < n) y[i] = fmaxf(a[i] + b[i], 0.f);
}

__global__ void relu_bwd_v4_kernel(const float4* __restrict__ y, const float4* __restrict__ dy, float4* __restrict__ dx, long long n4) {
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= n4) return;
  const float4 v = y[i], g = dy[i];
  dx[i] = make_float4(v.x > 0.f ? g.x : 0.f, v.y > 0.f ? g.y : 0.f, v.z > 0.f ? g.z : 0.f, v.w > 0.f ? g.w : 0.f);
}
__global__ void add_relu_fwd_v4_kernel(const float4* __restrict__ a, const float4* __restrict__ b, float4* __restrict__ y, long long n4) {
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= n4) return;
  const float4 u = a[i], v = b[i];
  y[i] = make_float4(fmaxf(u.x + v.x, 0.f), fmaxf(u.y + v.y, 0.f), fmaxf(u.z + v.z, 0.f), fmaxf(u.w + v.w, 0.f));
}
inline bool aligned16(const void* a, const void* b, const void* c) {
  return ((reinterpret_cast<uintptr_t>(a) | reinterpret_cast<uintptr_t>(b) | reinterpret_cast<uintptr_t>(c)) & 15) == 0;
}

// y[n,p,q,c] = max over the window; idx = r*k+s of the FIRST maximum in scan order (torch's tie rule).
__global__ void maxpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, uint8_t* __restrict__ idx, int N, int H, int W, int C,
                                   int k, int stride, int pad, int P, int Q) {
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  long long total = static_cast<long long>(N) * P * Q * C;
  if (i >= total) return;
  int c = i % C;
  long long t = i / C;
  int q = t % Q; t /= Q;
  int p = t % P;
  int n = static_cast<int>(t / P);
  float best = -INFINITY;
  int arg = -1;
  for (int r = 0; r < k; ++r) {
    int h = p * stride - pad + r;
    if (h < 0 || h >= H) continue;
    for (int s = 0; s < k; ++s) {
      int w = q * stride - pad + s;
      if (w < 0 || w >= W) continue;
      float v = x[((static_cast<long long>(n) * H + h) * W + w) * C + c];
      if (arg < 0 || v > best || v != v) { best = v; arg = r * k + s; }  // first max wins; NaN propagates
    }
  }
  y[i] = best;
  idx[i] = static_cast<uint8_t>(arg);
}

// dx[n,h,w,c] = sum of dy over the windows whose argmax is (h,w).  If relu_out != nullptr the ReLU
// that preceded the pool is folded in: windows whose pooled value is <= 0 pass no gradient.
__global__ void maxpool_bwd_kernel(const float* __restrict__ dy, const uint8_t* __restrict__ idx, const float* __restrict__ pooled,
                                   float* __restrict__ dx, int N, int H, int W, int C, int k, int stride, int pad, int P, int Q, int fuse_relu) {
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  long long total = static_cast<long long>(N) * H * W * C;
  if (i >= total) return;
  int c = i % C;
  long long t = i / C;
  int w = t % W; t /= W;
  int h = t % H;
  int n = static_cast<int>(t / H);
  float acc = 0.f;
  for (int r = 0; r < k; ++r) {
    int ph = h + pad - r;
    if (ph < 0 || ph % stride) continue;
    int p = ph / stride;
    if (p >= P) continue;
    for (int s = 0; s < k; ++s) {
      int qw = w + pad - s;
      if (qw < 0 || qw % stride) continue;
      int q = qw / stride;
      if (q >= Q) continue;
      long long o = ((static_cast<long long>(n) * P + p) * Q + q) * C + c;
      if (idx[o] == r * k + s && (!fuse_relu || pooled[o] > 0.f)) acc += dy[o];
    }
  }
  dx[i] = acc;
}

// Vector forms of the two pool kernels for C % 4 == 0 and < 2^31 elements: one thread per 4 channels,
// 32-bit index arithmetic (the scalar kernels above spend their time in 64-bit divisions).
__device__ __forceinline__ void take_max(float v, int code, float& best, int& arg) {
  if (arg < 0 || v > best || v != v) { best = v; arg = code; }
}
__global__ void __launch_bounds__(kThreads) maxpool_fwd_v4_kernel(const float4* __restrict__ x, float4* __restrict__ y, uchar4* __restrict__ idx,
                                                                  unsigned total4, int H, int W, int C4, int k, int stride, int pad, int P, int Q) {
  const unsigned i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= total4) return;
  const unsigned c = i % C4;
  unsigned t = i / C4;
  const int q = t % Q; t /= Q;
  const int p = t % P;
  const unsigned n = t / P;
  float4 best = make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
  int a0 = -1, a1 = -1, a2 = -1, a3 = -1;
  for (int r = 0; r < k; ++r) {
    const int h = p * stride - pad + r;
    if (h < 0 || h >= H) continue;
    for (int s = 0; s < k; ++s) {
      const int w = q * stride - pad + s;
      if (w < 0 || w >= W) continue;
      const float4 v = __ldg(x + ((n * H + h) * W + w) * C4 + c);
      const int code = r * k + s;
      take_max(v.x, code, best.x, a0); take_max(v.y, code, best.y, a1); take_max(v.z, code, best.z, a2); take_max(v.w, code, best.w, a3);
    }
  }
  y[i] = best;
  idx[i] = make_uchar4(a0, a1, a2, a3);
}
// kStride > 0 fixes the stride at compile time (the window search is otherwise bound by integer divisions)
template <int kStride>
__global__ void __launch_bounds__(kThreads) maxpool_bwd_v4_kernel(const float4* __restrict__ dy, const uchar4* __restrict__ idx,
                                                                  const float4* __restrict__ pooled, float4* __restrict__ dx, unsigned total4, int H,
                                                                  int W, int C4, int k, int stride_rt, int pad, int P, int Q, int fuse_relu) {
  const int stride = kStride > 0 ? kStride : stride_rt;
  const unsigned i = blockIdx.x * kThreads + threadIdx.x;
  if (i >= total4) return;
  const unsigned c = i % C4;
  unsigned t = i / C4;
  const int w = t % W; t /= W;
  const int h = t % H;
  const unsigned n = t / H;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int r = 0; r < k; ++r) {
    const int ph = h + pad - r;
    if (ph < 0 || ph % stride) continue;
    const int p = ph / stride;
    if (p >= P) continue;
    for (int s = 0; s < k; ++s) {
      const int qw = w + pad - s;
      if (qw < 0 || qw % stride) continue;
      const int q = qw / stride;
      if (q >= Q) continue;
      const unsigned o = ((n * P + p) * Q + q) * C4 + c;
      const uchar4 a = __ldg(idx + o);
      const float4 g = __ldg(dy + o);
      const int code = r * k + s;
      float4 pv = make_float4(1.f, 1.f, 1.f, 1.f);
      if (fuse_relu) pv = __ldg(pooled + o);
      acc.x += (a.x == code && pv.x > 0.f) ? g.x : 0.f; acc.y += (a.y == code && pv.y > 0.f) ? g.y : 0.f;
      acc.z += (a.z == code && pv.z > 0.f) ? g.z : 0.f; acc.w += (a.w == code && pv.w > 0.f) ? g.w : 0.f;
    }
  }
  dx[i] = acc;
}

__global__ void global_avgpool_fwd_kernel(const float* __restrict__ x, float* __restrict__ y, int N, int HW, int C) {
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= N * C) return;
  int c = i % C, n = i / C;
  float acc = 0.f;
  for (int j = 0; j < HW; ++j) acc += x[(static_cast<long long>(n) * HW + j) * C + c];
  y[i] = acc / HW;
}
__global__ void global_avgpool_bwd_kernel(const float* __restrict__ dy, float* __restrict__ dx, int N, int HW, int C) {
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= static_cast<long long>(N) * HW * C) return;
  int c = i % C;
  int n = static_cast<int>(i / (static_cast<long long>(HW) * C));
  dx[i] = dy[n * C + c] / HW;
}

// Layout change at the flatten boundary: the reference flattens NCHW (x.view(-1, 128*4*4),
// launch_dist.py:39) while activations here are NHWC.  to_nchw=1: y[n][c][hw] = x[n][hw][c].
__global__ void permute_nhwc_nchw_kernel(const float* __restrict__ x, float* __restrict__ y, int N, int HW, int C, int to_nchw) {
  long long i = blockIdx.x * static_cast<long long>(blockDim.x) + threadIdx.x;
  if (i >= static_cast<long long>(N) * HW * C) return;
  int per = HW * C;
  int n = static_cast<int>(i / per), j = static_cast<int>(i % per);
  if (to_nchw) { int c = j / HW, hw = j % HW; y[i] = x[static_cast<long long>(n) * per + hw * C + c]; }
  else { int hw = j / C, c = j % C; y[i] = x[static_cast<long long>(n) * per + c * HW + hw]; }
}

// ---- Linear (strict fp32) via the SIMT GEMM ----
struct RowMajorA { const float* a; int ld; __device__ float operator()(int m, int k) const { return a[static_cast<long long>(m) * ld + k]; } };
struct ColOfRowMajorB { const float* b; int ld; __device__ float operator()(int k, int n) const { return b[static_cast<long long>(n) * ld + k]; } };
struct RowMajorB { const float* b; int ld; __device__ float operator()(int k, int n) const { return b[static_cast<long long>(k) * ld + n]; } };
struct TransA { const float* a; int ld; __device__ float operator()(int m, int k) const { return a[static_cast<long long>(k) * ld + m]; } };
struct StoreBias {
  float* c; const float* bias; int ld;
  __device__ void operator()(int m, int n, float v, int) const { c[static_cast<long long>(m) * ld + n] = v + (bias ? bias[n] : 0.f); }
};
struct OnesColB {  // B(k, n) = x[k][n] for n < cols, 1 for the extra bias column
  const float* b; int ld; int cols;
  __device__ float operator()(int k, int n) const { return n < cols ? b[static_cast<long long>(k) * ld + n] : 1.0f; }
};
struct StoreWgrad {
  float* dw; float* db; int cols;
  __device__ void operator()(int m, int n, float v, int) const {
    if (n < cols) dw[static_cast<long long>(m) * cols + n] = v;
    else if (db) db[m] = v;
  }
};

// ---- CrossEntropy: one warp per row; last CTA reduces the per-row losses in index order ----
__device__ unsigned int g_ce_done_counter = 0;

__global__ void __launch_bounds__(kThreads) cross_entropy_kernel(const float* __restrict__ logits, const long long* __restrict__ target, int N,
                                                                 int C, float* __restrict__ row_loss, float* __restrict__ loss,
                                                                 float* __restrict__ dlogits, const float* __restrict__ grad_scale) {
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int row = blockIdx.x * (kThreads / 32) + warp;
  if (row < N) {
    const float* x = logits + static_cast<long long>(row) * C;
    float mx = -FLT_MAX;
    for (int c = lane; c < C; c += 32) mx = fmaxf(mx, x[c]);
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float sum = 0.f;
    for (int c = lane; c < C; c += 32) sum += expf(x[c] - mx);
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float lse = logf(sum) + mx;
    const int tgt = static_cast<int>(target[row]);
    if (lane == 0) row_loss[row] = lse - x[tgt];
    if (dlogits) {
      const float g = (grad_scale ? *grad_scale : 1.0f) / N;
      for (int c = lane; c < C; c += 32) dlogits[static_cast<long long>(row) * C + c] = (expf(x[c] - lse) - (c == tgt ? 1.f : 0.f)) * g;
    }
  }
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    unsigned int done = atomicAdd(&g_ce_done_counter, 1u);
    is_last = done == gridDim.x - 1;
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    __shared__ float part[kThreads];
    float acc = 0.f;
    // fixed partition + fixed tree => the same bits every run
    int per = (N + kThreads - 1) / kThreads, lo = threadIdx.x * per, hi = min(N, lo + per);
    for (int i = lo; i < hi; ++i) acc += __ldcg(row_loss + i);
    part[threadIdx.x] = acc;
    __syncthreads();
    for (int s = kThreads / 2; s; s >>= 1) {
      if (threadIdx.x < s) part[threadIdx.x] += part[threadIdx.x + s];
      __syncthreads();
    }
    if (threadIdx.x == 0) {
      *loss = part[0] / N;
      g_ce_done_counter = 0;
    }
  }
}

int launch_1d(long long n) { return pdl_ceil_div(n, kThreads); }

}  // namespace

PDL_API int pdl_relu_fwd(const float* x, float* y, long long n, cudaStream_t stream) {
  bool vec = ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(y)) & 15) == 0;
  long long n4 = vec ? n / 4 : 0;
  if (n4) relu_fwd_kernel<<<launch_1d(n4), kThreads, 0, stream>>>(reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(y), n4);
  if (n4 * 4 < n) relu_fwd_tail<<<launch_1d(n - n4 * 4), kThreads, 0, stream>>>(x, y, n4 * 4, n);
  PDL_LAUNCH_CHECK();
  return 0;
}
PDL_API int pdl_relu_bwd(const float* y, const float* dy, float* dx, long long n, cudaStream_t stream) {
  if (n && n % 4 == 0 && aligned16(y, dy, dx))
    relu_bwd_v4_kernel<<<launch_1d(n / 4), kThreads, 0, stream>>>(reinterpret_cast<const float4*>(y), reinterpret_cast<const float4*>(dy),
                                                                  reinterpret_cast<float4*>(dx), n / 4);
  else if (n) relu_bwd_kernel<<<launch_1d(n), kThreads, 0, stream>>>(y, dy, dx, n);
  PDL_LAUNCH_CHECK();
  return 0;
}
PDL_API int pdl_add_relu_fwd(const float* a, const float* b, float* y, long long n, cudaStream_t stream) {
  if (n && n % 4 == 0 && aligned16(a, b, y))
    add_relu_fwd_v4_kernel<<<launch_1d(n / 4), kThreads, 0, stream>>>(reinterpret_cast<const float4*>(a), reinterpret_cast<const float4*>(b),
                                                                      reinterpret_cast<float4*>(y), n / 4);
  else if (n) add_relu_fwd_kernel<<<launch_1d(n), kThreads, 0, stream>>>(a, b, y, n);
  PDL_LAUNCH_CHECK();
  return 0;
}
PDL_API int pdl_maxpool_fwd(const float* x, float* y, uint8_t* idx, int N, int H, int W, int C, int k, int stride, int pad, cudaStream_t stream) {
  int P = (H + 2 * pad - k) / stride + 1, Q = (W + 2 * pad - k) / stride + 1;
  long long total = static_cast<long long>(N) * P * Q * C;
  if (total && C % 4 == 0 && static_cast<long long>(N) * H * W * C < (1ll << 31)) {
    maxpool_fwd_v4_kernel<<<launch_1d(total / 4), kThreads, 0, stream>>>(reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(y),
                                                                         reinterpret_cast<uchar4*>(idx), static_cast<unsigned>(total / 4), H, W,
                                                                         C / 4, k, stride, pad, P, Q);
  } else if (total) maxpool_fwd_kernel<<<launch_1d(total), kThreads, 0, stream>>>(x, y, idx, N, H, W, C, k, stride, pad, P, Q);
  PDL_LAUNCH_CHECK();
  return 0;
}
PDL_API int pdl_maxpool_bwd(const float* dy, const uint8_t* idx, const float* pooled, float* dx, int N, int H, int W, int C, int k, int stride,
                            int pad, int fuse_relu, cudaStream_t stream) {
  int P = (H + 2 * pad - k) / stride + 1, Q = (W + 2 * pad - k) / stride + 1;
  long long total = static_cast<long long>(N) * H * W * C;
  if (total && C % 4 == 0 && total < (1ll << 31)) {
    auto kern = stride == 1 ? maxpool_bwd_v4_kernel<1> : stride == 2 ? maxpool_bwd_v4_kernel<2> : maxpool_bwd_v4_kernel<0>;
    kern<<<launch_1d(total / 4), kThreads, 0, stream>>>(
        reinterpret_cast<const float4*>(dy), reinterpret_cast<const uchar4*>(idx), reinterpret_cast<const float4*>(pooled),
        reinterpret_cast<float4*>(dx), static_cast<unsigned>(total / 4), H, W, C / 4, k, stride, pad, P, Q, fuse_relu);
  } else if (total) maxpool_bwd_kernel<<<launch_1d(total), kThreads, 0, stream>>>(dy, idx, pooled, dx, N, H, W, C, k, stride, pad, P, Q, fuse_relu);
  PDL_LAUNCH_CHECK();
  return 0;
}
PDL_API int pdl_global_avgpool_fwd(const float* x, float* y, int N, int HW, int C, cudaStream_t stream) {
  global_avgpool_fwd_kernel<<<launch_1d(static_cast<long long>(N) * C), kThreads, 0, stream>>>(x, y, N, HW, C);
  PDL_LAUNCH_CHECK();
  return 0;
}
PDL_API int pdl_global_avgpool_bwd(const float* dy, float* dx, int N, int HW, int C, cudaStream_t stream) {
  global_avgpool_bwd_kernel<<<launch_1d(static_cast<long long>(N) * HW * C), kThreads, 0, stream>>>(dy, dx, N, HW, C);
  PDL_LAUNCH_CHECK();
  return 0;
}

namespace {
// part[s][c] = sum of x[r][c] over the rows r of split s (NHWC rows [M][C]): the bias gradient of a convolution as partial
// sums that the weight-gradient kernel's deterministic second pass folds in (reference: the db reduction autograd runs for
// nn.Conv2d(bias=True), launch_dist.py:15-25).  256 threads = 8 row lanes x 32 channels; fixed summation order.
__global__ void __launch_bounds__(256) colsum_partial_kernel(const float* __restrict__ x, float* __restrict__ part, long long M, int C,
                                                             long long rows_per_split) {
  const int cl = threadIdx.x % 32, rl = threadIdx.x / 32;
  const int c = blockIdx.x * 32 + cl;
  const long long r_lo = blockIdx.y * rows_per_split, r_hi = min(M, r_lo + rows_per_split);
  float s = 0.f;
  if (c < C) {
#pragma unroll 4
    for (long long r = r_lo + rl; r < r_hi; r += 8) s += __ldg(x + r * C + c);
  }
  __shared__ float sm[8][33];
  sm[rl][cl] = s;
  __syncthreads();
  if (rl == 0 && c < C) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += sm[i][cl];
    part[static_cast<long long>(blockIdx.y) * C + c] = t;
  }
}
}  // namespace

// Number of row splits pdl_colsum_partial uses (= rows of `part`).
PDL_API int pdl_colsum_splits(long long M, int C) {
  long long want = 296 / ((C + 31) / 32);  // about two waves of CTAs in total
  if (want < 1) want = 1;
  long long by_rows = (M + 63) / 64;       // at least 64 rows per split
  if (want > by_rows) want = by_rows;
  return static_cast<int>(want < 1 ? 1 : want);
}
PDL_API int pdl_colsum_partial(const float* x, float* part, long long M, int C, cudaStream_t stream) {
  const int splits = pdl_colsum_splits(M, C);
  const long long rps = (M + splits - 1) / splits;
  colsum_partial_kernel<<<dim3((C + 31) / 32, splits), 256, 0, stream>>>(x, part, M, C, rps);
  PDL_LAUNCH_CHECK();
  return 0;
}

PDL_API int pdl_permute_nhwc_nchw(const float* x, float* y, int N, int HW, int C, int to_nchw, cudaStream_t stream) {
  long long total = static_cast<long long>(N) * HW * C;
  if (total) permute_nhwc_nchw_kernel<<<launch_1d(total), kThreads, 0, stream>>>(x, y, N, HW, C, to_nchw);
  PDL_LAUNCH_CHECK();
  return 0;
}

namespace {
// Linear forward for a batch of at most a few hundred rows and a long K (the classifier heads: 2048 -> 1000 at batch
// 64).  The generic 64x64 tile kernel gives such a problem ~16 CTAs that each walk K serially with strided weight reads
// (237 us on a B200); here a warp owns ONE output feature for a block of kLinRows rows: lanes stride K with float4
// loads (the weight row is read once, coalesced; the activation rows come from L1/L2), kLinRows accumulators per lane,
// fixed-order shuffle reduction at the end.  Strict fp32, deterministic.  K % 128 == 0.
constexpr int kLinRows = 32;
__global__ void __launch_bounds__(256) linear_fwd_rows_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ b,
                                                              float* __restrict__ y, int M, int N, int K) {
  const int lane = threadIdx.x % 32, n = blockIdx.x * 8 + threadIdx.x / 32, m0 = blockIdx.y * kLinRows;
  if (n >= N) return;
  float acc[kLinRows];
#pragma unroll
  for (int r = 0; r < kLinRows; ++r) acc[r] = 0.f;
  const float4* wr = reinterpret_cast<const float4*>(w + static_cast<long long>(n) * K);
  const int rows = min(kLinRows, M - m0);
  for (int k4 = lane; k4 < K / 4; k4 += 32) {
    const float4 wv = __ldg(wr + k4);
#pragma unroll
    for (int r = 0; r < kLinRows; ++r) {
      if (r < rows) {
        const float4 xv = __ldg(reinterpret_cast<const float4*>(x + static_cast<long long>(m0 + r) * K) + k4);
        acc[r] = fmaf(xv.x, wv.x, fmaf(xv.y, wv.y, fmaf(xv.z, wv.z, fmaf(xv.w, wv.w, acc[r]))));
      }
    }
  }
  const float bias = b ? b[n] : 0.f;
#pragma unroll
  for (int r = 0; r < kLinRows; ++r) {
    float v = acc[r];
#pragma unroll
    for (int o = 16; o; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    if (lane == 0 && r < rows) y[static_cast<long long>(m0 + r) * N + n] = v + bias;
  }
}
}  // namespace

// y[M,N] = x[M,K] * w[N,K]^T + b
PDL_API int pdl_linear_fwd(const float* x, const float* w, const float* b, float* y, int M, int N, int K, cudaStream_t stream) {
  if (M > 0 && M <= 512 && K % 128 == 0 && K >= 512 && N >= 64 &&
      ((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(w)) & 15) == 0) {
    linear_fwd_rows_kernel<<<dim3((N + 7) / 8, (M + kLinRows - 1) / kLinRows), 256, 0, stream>>>(x, w, b, y, M, N, K);
    PDL_LAUNCH_CHECK();
    return 0;
  }
  return simt::launch(M, N, K, 1, RowMajorA{x, K}, ColOfRowMajorB{w, K}, StoreBias{y, b, N}, stream);
}
// dx[M,K] = dy[M,N] * w[N,K]
PDL_API int pdl_linear_dgrad(const float* dy, const float* w, float* dx, int M, int N, int K, cudaStream_t stream) {
  return simt::launch(M, K, N, 1, RowMajorA{dy, N}, RowMajorB{w, K}, StoreBias{dx, nullptr, K}, stream);
}
// dw[N,K] = dy[M,N]^T * x[M,K];  db[N] = column sums of dy (folded in as one extra GEMM column)
PDL_API int pdl_linear_wgrad(const float* x, const float* dy, float* dw, float* db, int M, int N, int K, cudaStream_t stream) {
  return simt::launch(N, K + (db ? 1 : 0), M, 1, TransA{dy, N}, OnesColB{x, K, K}, StoreWgrad{dw, db, K}, stream);
}

// loss (scalar, mean over rows) and, if dlogits != nullptr, d loss / d logits * (*grad_scale or 1).
PDL_API int pdl_cross_entropy(const float* logits, const long long* target, int N, int C, float* row_loss, float* loss, float* dlogits,
                              const float* grad_scale, cudaStream_t stream) {
  int blocks = pdl_ceil_div(N, kThreads / 32);
  cross_entropy_kernel<<<blocks, kThreads, 0, stream>>>(logits, target, N, C, row_loss, loss, dlogits, grad_scale);
  PDL_LAUNCH_CHECK();
  return 0;
}

// Running training metrics kept on the device (reference example_launch.py:91-93 does argmax / eq / sum / .item() and
// loss.item() -- two host syncs -- every step; here the step only accumulates and the host reads when it prints).
// acc[0] += *loss (if loss), acc[1] += #rows whose arg-max (first maximum) equals the label, acc[2] += N.
// One warp per row; counts are integers carried in doubles, so the atomic adds commute exactly.
// (v, c) replaces (best, arg)?  Larger value wins, a NaN counts as the maximum (as in torch), ties go to the lower index;
// index 0x7fffffff marks "nothing seen yet".
__device__ __forceinline__ bool argmax_beats(float v, int c, float best, int arg) {
  if (c == 0x7fffffff) return false;
  if (arg == 0x7fffffff) return true;
  bool nan_v = v != v, nan_b = best != best;
  if (nan_v != nan_b) return nan_v;
  if (nan_v || v == best) return c < arg;
  return v > best;
}

__global__ void __launch_bounds__(kThreads) step_metrics_kernel(const float* __restrict__ logits, const long long* __restrict__ target,
                                                                const float* __restrict__ loss, int N, int C, double* __restrict__ acc) {
  __shared__ int hits[kThreads / 32];
  int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int row = blockIdx.x * (kThreads / 32) + warp;
  int hit = 0;
  if (row < N) {
    const float* r = logits + (size_t)row * C;
    float best = -INFINITY;
    int arg = 0x7fffffff;
    for (int c = lane; c < C; c += 32) {
      float v = r[c];
      if (argmax_beats(v, c, best, arg)) { best = v; arg = c; }
    }
    for (int o = 16; o; o >>= 1) {
      float ob = __shfl_xor_sync(0xffffffffu, best, o);
      int oa = __shfl_xor_sync(0xffffffffu, arg, o);
      if (argmax_beats(ob, oa, best, arg)) { best = ob; arg = oa; }
    }
    hit = (long long)arg == target[row];
  }
  if (lane == 0) hits[warp] = hit;
  __syncthreads();
  if (threadIdx.x == 0) {
    int h = 0;
    for (int w = 0; w < kThreads / 32; ++w) h += hits[w];
    if (h) atomicAdd(&acc[1], (double)h);
    if (blockIdx.x == 0) {
      if (loss) atomicAdd(&acc[0], (double)*loss);
      atomicAdd(&acc[2], (double)N);
    }
  }
}

// Refresh a captured graph's static inputs (images + labels) with ONE launch instead of two device-to-device memcpys:
// at 0.12 ms per step every stream operation in front of the graph (~2 us each) is visible (reference loop:
// launch_dist.py:82-83 issues the two copies).  16-byte vector copies where both pointers are aligned; tails and
// misaligned ranges go byte by byte.
__global__ void __launch_bounds__(kThreads) copy_pair_kernel(char* __restrict__ d0, const char* __restrict__ s0, size_t n0,
                                                             char* __restrict__ d1, const char* __restrict__ s1, size_t n1) {
  size_t tid = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
  for (int which = 0; which < 2; ++which) {
    char* d = which ? d1 : d0;
    const char* s = which ? s1 : s0;
    size_t n = which ? n1 : n0;
    bool vec = ((reinterpret_cast<uintptr_t>(d) | reinterpret_cast<uintptr_t>(s)) & 15) == 0;
    size_t n16 = vec ? n / 16 : 0;
    for (size_t i = tid; i < n16; i += stride) reinterpret_cast<float4*>(d)[i] = reinterpret_cast<const float4*>(s)[i];
    for (size_t i = n16 * 16 + tid; i < n; i += stride) d[i] = s[i];
  }
}

PDL_API int pdl_copy_pair(void* d0, const void* s0, long long n0, void* d1, const void* s1, long long n1, cudaStream_t stream) {
  if (n0 < 0 || n1 < 0) return pdl_set_error("copy_pair: negative size"), -1;
  if (n0 + n1 == 0) return 0;
  long long work = (n0 + n1) / 16 + 1;
  int blocks = (int)std::min<long long>(pdl_ceil_div(work, (long long)kThreads), 148 * 8);
  copy_pair_kernel<<<blocks, kThreads, 0, stream>>>((char*)d0, (const char*)s0, (size_t)n0, (char*)d1, (const char*)s1, (size_t)n1);
  PDL_LAUNCH_CHECK();
  return 0;
}

PDL_API int pdl_step_metrics(const float* logits, const long long* target, const float* loss, int N, int C, double* acc, cudaStream_t stream) {
  if (N <= 0 || C <= 0) return 0;
  step_metrics_kernel<<<pdl_ceil_div(N, kThreads / 32), kThreads, 0, stream>>>(logits, target, loss, N, C, acc);
  PDL_LAUNCH_CHECK();
  return 0;
}
