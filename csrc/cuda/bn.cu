// BatchNorm2d for the ResNets the reference trains (torchvision resnet18 at example_launch.py:26; ResNet-50
// per BASELINE.json), NHWC fp32, with the block's ReLU and residual add fused in:
//     y = act( (x - mean) * invstd * gamma + beta  [+ residual] ),   act = ReLU or identity
// Training mode: per-channel batch statistics in one pass over x (per-thread shifted sums turned into
// Welford triples and merged in a fixed order => deterministic), running-stat update with the unbiased
// variance exactly like torch.  Backward: dgamma/dbeta, dx in the standard closed form and (for the residual
// form) the gradient of the skip connection, with the ReLU mask recomputed from x (bit-exact: same fmaf,
// same saved scale/shift) or taken from y when a residual was added.
//   x, y, dy, dx, residual : [M][C] with M = N*H*W rows, channels contiguous, C % 4 == 0
//   saved                  : [4][C] = mean, invstd, scale, shift (written by forward, read by backward)
// Everything is HBM-bound: float4 accesses, >= 8 CTAs per SM worth of splits, no 64-bit division.
#include <type_traits>

#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kCG = 32;              // channels per CTA of the statistics kernels (one 128-byte row segment)
constexpr int kQL = kCG / 4;         // float4 lanes across the channel group
constexpr int kRL = kThreads / kQL;  // row lanes per CTA

struct Welford { float n, mean, m2; };
__device__ __forceinline__ Welford merge(Welford a, Welford b) {
  if (b.n == 0.f) return a;
  if (a.n == 0.f) return b;
  const float n = a.n + b.n, d = b.mean - a.mean;
  return {n, a.mean + d * (b.n / n), a.m2 + b.m2 + d * d * (a.n * b.n / n)};
}
__device__ __forceinline__ Welford from_shifted(float n, float pivot, float s, float q) {
  if (n == 0.f) return {0.f, 0.f, 0.f};
  const float m = s / n;
  return {n, pivot + m, fmaxf(q - s * m, 0.f)};
}

// partial[(split * C + c) * 3 + {0,1,2}] = (count, mean, M2) of x[rows of this split][c]
struct FinalizeArgs {
  const float* gamma; const float* beta; float* running_mean; float* running_var; float momentum, eps; float* saved;
  long long* num_batches_tracked;
};
__device__ void finalize_group(const float* partial, int splits, int C, int cgroup, const FinalizeArgs& f);
// Per channel group: CTAs of the running statistics kernel that have published their partials.  ONE set of counters per
// device: BatchNorm launches (forward statistics, backward) must be stream-ordered with respect to each other -- two of them
// overlapping on different streams or graph branches would corrupt each other's last-CTA detection.  This package issues all
// of them on the compute stream; the rule is repeated in include/pdl_b200.h and INTEGRATION.md.
__device__ unsigned int g_bn_done[2048];

// Returns true in the LAST CTA of this channel group to finish (its reads of the other CTAs' partials are ordered
// after their writes); resets the counter for the next launch.  Launches are stream-ordered, so one array suffices.
__device__ __forceinline__ bool last_cta_of_group() {
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const unsigned done = atomicAdd(&g_bn_done[blockIdx.x], 1u);
    is_last = done == gridDim.y - 1;
    if (is_last) g_bn_done[blockIdx.x] = 0;
    __threadfence();
  }
  __syncthreads();
  return is_last;
}

__global__ void __launch_bounds__(kThreads) bn_stats_kernel(const float* __restrict__ x, float* __restrict__ partial, long long M, int C,
                                                            long long rows_per_split, const FinalizeArgs fin) {
  const int ql = threadIdx.x % kQL, rl = threadIdx.x / kQL;
  const int c = blockIdx.x * kCG + ql * 4;
  const long long r_lo = blockIdx.y * rows_per_split, r_hi = min(M, r_lo + rows_per_split);
  __shared__ float shs[kRL][kCG + 1], shq[kRL][kCG + 1], shn[kRL], shk[kCG];
  if (c < C) {
    const float4* xp = reinterpret_cast<const float4*>(x + c);
    const long long ld = C / 4;
    // pivot: first row of the split (inside the data range, shared by all row lanes); a trailing split may be empty
    const float4 k = r_lo < M ? __ldg(xp + r_lo * ld) : make_float4(0.f, 0.f, 0.f, 0.f);
    float4 s = make_float4(0.f, 0.f, 0.f, 0.f), q = s;
    float n = 0.f;
#pragma unroll 8
    for (long long r = r_lo + rl; r < r_hi; r += kRL) {
      const float4 v = __ldg(xp + r * ld);
      const float a = v.x - k.x, b = v.y - k.y, cc = v.z - k.z, d = v.w - k.w;
      s.x += a; s.y += b; s.z += cc; s.w += d;
      q.x = fmaf(a, a, q.x); q.y = fmaf(b, b, q.y); q.z = fmaf(cc, cc, q.z); q.w = fmaf(d, d, q.w);
      n += 1.f;
    }
    shs[rl][ql * 4 + 0] = s.x; shs[rl][ql * 4 + 1] = s.y; shs[rl][ql * 4 + 2] = s.z; shs[rl][ql * 4 + 3] = s.w;
    shq[rl][ql * 4 + 0] = q.x; shq[rl][ql * 4 + 1] = q.y; shq[rl][ql * 4 + 2] = q.z; shq[rl][ql * 4 + 3] = q.w;
    if (ql == 0) shn[rl] = n;
    if (rl == 0) { shk[ql * 4 + 0] = k.x; shk[ql * 4 + 1] = k.y; shk[ql * 4 + 2] = k.z; shk[ql * 4 + 3] = k.w; }
  }
  __syncthreads();
  if (threadIdx.x < kCG && blockIdx.x * kCG + threadIdx.x < C) {  // one warp: fixed-order sums over the row lanes
    const int ch = threadIdx.x;
    float n = 0.f, ss = 0.f, qq = 0.f;
#pragma unroll
    for (int i = 0; i < kRL; ++i) { n += shn[i]; ss += shs[i][ch]; qq += shq[i][ch]; }
    const Welford t = from_shifted(n, shk[ch], ss, qq);
    float* o = partial + (static_cast<long long>(blockIdx.y) * C + blockIdx.x * kCG + ch) * 3;
    o[0] = t.n; o[1] = t.mean; o[2] = t.m2;
  }
  if (last_cta_of_group()) finalize_group(partial, gridDim.y, C, blockIdx.x, fin);
}

// Merge the splits of one channel group without per-split divisions: mean = sum(n_i mean_i) / n, then
// M2 = sum(M2_i + n_i (mean_i - mean)^2).  Runs in the last CTA of the group: 32 channels x 8 slices of the split
// list, fixed order; partials come from other CTAs, so they are read through L2 (__ldcg).
constexpr int kFinSlices = kThreads / kCG;
__device__ void finalize_group(const float* partial, int splits, int C, int cgroup, const FinalizeArgs& f) {
  const float* gamma = f.gamma; const float* beta = f.beta;
  float* running_mean = f.running_mean; float* running_var = f.running_var; float* saved = f.saved;
  const float momentum = f.momentum, eps = f.eps;
  if (f.num_batches_tracked && cgroup == 0 && threadIdx.x == 0) *f.num_batches_tracked += 1;  // nn.BatchNorm2d's step counter
  const int ch = threadIdx.x % kCG, sl = threadIdx.x / kCG;
  const int c = cgroup * kCG + ch;
  __shared__ float sh_a[kFinSlices][kCG + 1], sh_b[kFinSlices][kCG + 1];
  float sn = 0.f, sm = 0.f;
  if (c < C) {
#pragma unroll 8
    for (int s = sl; s < splits; s += kFinSlices) {
      const float* p = partial + (static_cast<long long>(s) * C + c) * 3;
      const float pn = __ldcg(p);
      sn += pn;
      sm = fmaf(pn, __ldcg(p + 1), sm);
    }
  }
  sh_a[sl][ch] = sn; sh_b[sl][ch] = sm;
  __syncthreads();
  float n = 0.f, m = 0.f;
#pragma unroll
  for (int i = 0; i < kFinSlices; ++i) { n += sh_a[i][ch]; m += sh_b[i][ch]; }
  const float mean = m / n;
  __syncthreads();
  float m2 = 0.f;
  if (c < C) {
#pragma unroll 8
    for (int s = sl; s < splits; s += kFinSlices) {
      const float* p = partial + (static_cast<long long>(s) * C + c) * 3;
      const float d = __ldcg(p + 1) - mean;
      m2 += fmaf(__ldcg(p) * d, d, __ldcg(p + 2));
    }
  }
  sh_a[sl][ch] = m2;
  __syncthreads();
  if (sl != 0 || c >= C) return;
  m2 = 0.f;
#pragma unroll
  for (int i = 0; i < kFinSlices; ++i) m2 += sh_a[i][ch];
  const Welford t{n, mean, m2};
  const float var = t.m2 / t.n;  // biased: used for normalisation
  const float invstd = rsqrtf(var + eps);
  const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  const float scale = g * invstd;
  saved[c] = t.mean;
  saved[C + c] = invstd;
  saved[2 * C + c] = scale;
  saved[3 * C + c] = fmaf(-t.mean, scale, b);
  if (running_mean) {
    const float unbiased = t.n > 1.f ? t.m2 / (t.n - 1.f) : var;
    running_mean[c] = (1.f - momentum) * running_mean[c] + momentum * t.mean;
    running_var[c] = (1.f - momentum) * running_var[c] + momentum * unbiased;
  }
}

__global__ void bn_eval_coeffs_kernel(int C, const float* __restrict__ gamma, const float* __restrict__ beta, const float* __restrict__ rm,
                                      const float* __restrict__ rv, float eps, float* __restrict__ scale, float* __restrict__ shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  const float invstd = rsqrtf(rv[c] + eps), g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  scale[c] = g * invstd;
  shift[c] = fmaf(-rm[c], g * invstd, b);
}

__device__ __forceinline__ int quad_channel(long long i, int C4) {
  // 32-bit modulo whenever the index fits (64-bit division is ~10x slower and would dominate the kernel)
  return 4 * ((i >> 32) == 0 ? static_cast<int>(static_cast<unsigned>(i) % static_cast<unsigned>(C4)) : static_cast<int>(i % C4));
}

// y = act(x * scale[c] + shift[c] [+ res])
template <bool kRelu, bool kRes>
__global__ void __launch_bounds__(kThreads) bn_apply_kernel(const float4* __restrict__ x, const float4* __restrict__ res, float4* __restrict__ y,
                                                            const float* __restrict__ scale, const float* __restrict__ shift, long long total4,
                                                            int C4, uint8_t* __restrict__ relu_mask) {
  const long long i = blockIdx.x * static_cast<long long>(kThreads) + threadIdx.x;
  if (i >= total4) return;
  const int c = quad_channel(i, C4);
  const float4 v = x[i], s = *reinterpret_cast<const float4*>(scale + c), h = *reinterpret_cast<const float4*>(shift + c);
  float4 o = make_float4(fmaf(v.x, s.x, h.x), fmaf(v.y, s.y, h.y), fmaf(v.z, s.z, h.z), fmaf(v.w, s.w, h.w));
  if (kRes) { const float4 r = res[i]; o.x += r.x; o.y += r.y; o.z += r.z; o.w += r.w; }
  if (kRelu && kRes && relu_mask)  // one byte per 4 elements: what backward needs of y, at 1/16 of its size
    relu_mask[i] = static_cast<uint8_t>((o.x > 0.f) | ((o.y > 0.f) << 1) | ((o.z > 0.f) << 2) | ((o.w > 0.f) << 3));
  if (kRelu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
  y[i] = o;
}

// The gradient that reaches the normalisation: dy through the activation.  kMask 0: no activation; 1: ReLU mask
// recomputed from x (no residual); 2: ReLU mask bits `m` stored by the forward pass (residual form).
template <int kMask>
__device__ __forceinline__ float4 act_grad(float4 g, float4 v, unsigned m, float4 s, float4 h) {
  if (kMask == 1) {
    g.x = fmaf(v.x, s.x, h.x) > 0.f ? g.x : 0.f; g.y = fmaf(v.y, s.y, h.y) > 0.f ? g.y : 0.f;
    g.z = fmaf(v.z, s.z, h.z) > 0.f ? g.z : 0.f; g.w = fmaf(v.w, s.w, h.w) > 0.f ? g.w : 0.f;
  } else if (kMask == 2) {
    g.x = (m & 1u) ? g.x : 0.f; g.y = (m & 2u) ? g.y : 0.f; g.z = (m & 4u) ? g.z : 0.f; g.w = (m & 8u) ? g.w : 0.f;
  }
  return g;
}

struct BwdFinalizeArgs { float M; const float* gamma; float* dgamma; float* dbeta; float* coef; };
__device__ void bwd_finalize_group(const float* partial, int splits, int C, int cgroup, const float* saved, const BwdFinalizeArgs& f);

// partial[(split * C + c) * 2 + {0,1}] = (sum g, sum g * xhat) over the rows of the split
template <int kMask>
__global__ void __launch_bounds__(kThreads) bn_bwd_stats_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                                const float* __restrict__ dy2,  // second addend of the incoming gradient or nullptr
                                                                const uint8_t* __restrict__ relu_mask, const float* __restrict__ saved,
                                                                float* __restrict__ partial, long long M, int C, long long rows_per_split,
                                                                const BwdFinalizeArgs fin) {
  const int ql = threadIdx.x % kQL, rl = threadIdx.x / kQL;
  const int c = blockIdx.x * kCG + ql * 4;
  const long long r_lo = blockIdx.y * rows_per_split, r_hi = min(M, r_lo + rows_per_split);
  __shared__ float a1[kRL][kCG + 1], a2[kRL][kCG + 1];
  if (c < C) {
    const long long ld = C / 4;
    const float4* xp = reinterpret_cast<const float4*>(x + c);
    const float4* gp = reinterpret_cast<const float4*>(dy + c);
    const float4* gp2 = dy2 ? reinterpret_cast<const float4*>(dy2 + c) : nullptr;
    const uint8_t* mp = relu_mask + c / 4;
    const float4 mu = *reinterpret_cast<const float4*>(saved + c), is = *reinterpret_cast<const float4*>(saved + C + c);
    const float4 sc = *reinterpret_cast<const float4*>(saved + 2 * C + c), sf = *reinterpret_cast<const float4*>(saved + 3 * C + c);
    float4 s1 = make_float4(0.f, 0.f, 0.f, 0.f), s2 = s1;
#pragma unroll 4
    for (long long r = r_lo + rl; r < r_hi; r += kRL) {
      const float4 v = __ldg(xp + r * ld);
      float4 g = __ldg(gp + r * ld);
      if (gp2) { const float4 g2 = __ldg(gp2 + r * ld); g.x += g2.x; g.y += g2.y; g.z += g2.z; g.w += g2.w; }
      unsigned m = 0;
      if (kMask == 2) m = __ldg(mp + r * ld);
      g = act_grad<kMask>(g, v, m, sc, sf);
      s1.x += g.x; s1.y += g.y; s1.z += g.z; s1.w += g.w;
      s2.x = fmaf(g.x, (v.x - mu.x) * is.x, s2.x); s2.y = fmaf(g.y, (v.y - mu.y) * is.y, s2.y);
      s2.z = fmaf(g.z, (v.z - mu.z) * is.z, s2.z); s2.w = fmaf(g.w, (v.w - mu.w) * is.w, s2.w);
    }
    a1[rl][ql * 4 + 0] = s1.x; a1[rl][ql * 4 + 1] = s1.y; a1[rl][ql * 4 + 2] = s1.z; a1[rl][ql * 4 + 3] = s1.w;
    a2[rl][ql * 4 + 0] = s2.x; a2[rl][ql * 4 + 1] = s2.y; a2[rl][ql * 4 + 2] = s2.z; a2[rl][ql * 4 + 3] = s2.w;
  }
  __syncthreads();
  const int ch = threadIdx.x % kCG, grp = threadIdx.x / kCG;
  if (grp < 2 && blockIdx.x * kCG + ch < C) {  // warp 0 sums a1, warp 1 sums a2, both in row-lane order
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < kRL; ++i) t += grp == 0 ? a1[i][ch] : a2[i][ch];
    partial[(static_cast<long long>(blockIdx.y) * C + blockIdx.x * kCG + ch) * 2 + grp] = t;
  }
  if (last_cta_of_group()) bwd_finalize_group(partial, gridDim.y, C, blockIdx.x, saved, fin);
}

// dgamma, dbeta and the per-channel coefficients of dx = a*g + b*x + d (last CTA of the channel group)
__device__ void bwd_finalize_group(const float* partial, int splits, int C, int cgroup, const float* saved, const BwdFinalizeArgs& f) {
  const float M = f.M; const float* gamma = f.gamma; float* dgamma = f.dgamma; float* dbeta = f.dbeta; float* coef = f.coef;
  const int ch = threadIdx.x % kCG, sl = threadIdx.x / kCG;
  const int c = cgroup * kCG + ch;
  __shared__ float h1[kFinSlices][kCG + 1], h2[kFinSlices][kCG + 1];
  float s1 = 0.f, s2 = 0.f;
  if (c < C) {
#pragma unroll 8
    for (int s = sl; s < splits; s += kFinSlices) {
      const float2 p = __ldcg(reinterpret_cast<const float2*>(partial + (static_cast<long long>(s) * C + c) * 2));
      s1 += p.x; s2 += p.y;
    }
  }
  h1[sl][ch] = s1; h2[sl][ch] = s2;
  __syncthreads();
  if (sl != 0 || c >= C) return;
#pragma unroll
  for (int i = 1; i < kFinSlices; ++i) { s1 += h1[i][ch]; s2 += h2[i][ch]; }
  if (dgamma) dgamma[c] = s2;
  if (dbeta) dbeta[c] = s1;
  // dx = g*is * (dy - s1/M - xhat * s2/M),  xhat = (x - mu) * is
  const float g = gamma ? gamma[c] : 1.f, mu = saved[c], is = saved[C + c];
  const float a = g * is, k = a * is * s2 / M;
  coef[c] = a;
  coef[C + c] = -k;
  coef[2 * C + c] = -a * s1 / M + k * mu;
}

template <int kMask, bool kRes>
__global__ void __launch_bounds__(kThreads) bn_bwd_apply_kernel(const float4* __restrict__ x, const float4* __restrict__ dy,
                                                                const float4* __restrict__ dy2,
                                                                const uint8_t* __restrict__ relu_mask, float4* __restrict__ dx,
                                                                float4* __restrict__ dres, const float* __restrict__ coef,
                                                                const float* __restrict__ saved, long long total4, int C4) {
  const long long i = blockIdx.x * static_cast<long long>(kThreads) + threadIdx.x;
  if (i >= total4) return;
  const int c = quad_channel(i, C4), C = 4 * C4;
  const float4 v = x[i];
  float4 g = dy[i];
  if (dy2) { const float4 g2 = dy2[i]; g.x += g2.x; g.y += g2.y; g.z += g2.z; g.w += g2.w; }
  unsigned m = 0;
  if (kMask == 2) m = relu_mask[i];
  if (kMask != 0) {
    const float4 sc = *reinterpret_cast<const float4*>(saved + 2 * C + c), sf = *reinterpret_cast<const float4*>(saved + 3 * C + c);
    g = act_grad<kMask>(g, v, m, sc, sf);
  }
  const float4 a = *reinterpret_cast<const float4*>(coef + c), b = *reinterpret_cast<const float4*>(coef + C + c),
               d = *reinterpret_cast<const float4*>(coef + 2 * C + c);
  dx[i] = make_float4(fmaf(a.x, g.x, fmaf(b.x, v.x, d.x)), fmaf(a.y, g.y, fmaf(b.y, v.y, d.y)), fmaf(a.z, g.z, fmaf(b.z, v.z, d.z)),
                      fmaf(a.w, g.w, fmaf(b.w, v.w, d.w)));
  if (kRes) dres[i] = g;
}

int choose_splits(long long M, int C) {
  const int cgroups = (C + kCG - 1) / kCG;
  // Between one wave (148 CTAs) and 4 CTAs per SM, sized so that a CTA streams >= ~64 KB: the merge in each channel
  // group's last CTA walks the split list (8 slices), and with 296 splits of 20 KB each that tail was HALF of the
  // kernel on ResNet-50's 13 MB layers (ncu: 14.7 us at 10 % DRAM utilisation).
  long long ctas = M * C * 4 / 65536;
  if (ctas < 148) ctas = 148;
  if (ctas > 4 * 148) ctas = 4 * 148;
  long long s = (ctas + cgroups - 1) / cgroups;
  const long long by_rows = (M + 127) / 128;  // at least 4 rows per thread
  if (s > by_rows) s = by_rows;
  if (s < 1) s = 1;
  if (s > 296) s = 296;
  return static_cast<int>(s);
}

int check_c(int C) {
  if (C % 4) { pdl_set_error("batch norm: channel count %d is not a multiple of 4", C); return -1; }
  return 0;
}

template <typename F>
void launch_apply(F&& f, int relu, bool res) {
  if (relu) { if (res) f(std::true_type{}, std::true_type{}); else f(std::true_type{}, std::false_type{}); }
  else      { if (res) f(std::false_type{}, std::true_type{}); else f(std::false_type{}, std::false_type{}); }
}

void apply(const float* x, const float* residual, float* y, const float* scale, const float* shift, long long M, int C, int relu,
           uint8_t* relu_mask, cudaStream_t stream) {
  const long long total4 = M * C / 4;
  const unsigned blocks = static_cast<unsigned>(pdl_ceil_div(total4, kThreads));
  launch_apply([&](auto r, auto s) {
    bn_apply_kernel<decltype(r)::value, decltype(s)::value><<<blocks, kThreads, 0, stream>>>(
        reinterpret_cast<const float4*>(x), reinterpret_cast<const float4*>(residual), reinterpret_cast<float4*>(y), scale, shift, total4, C / 4,
        relu_mask);
  }, relu, residual != nullptr);
}

}  // namespace

// floats of scratch the BN entry points need: split partials + 3 per-channel coefficient rows
PDL_API long long pdl_bn_scratch_floats(long long M, int C) { return static_cast<long long>(choose_splits(M, C)) * C * 3 + 3ll * C; }

// Training forward.  num_batches_tracked (may be NULL) is incremented on the device.  saved [4][C] (mean, invstd,
// scale, shift) is kept for backward.  relu != 0 fuses the activation, residual != nullptr adds the skip connection
// before it; with both, relu_mask (M*C/4 bytes, 4 sign bits each) receives what backward needs of the output.
PDL_API int pdl_bn_fwd_train(const float* x, const float* residual, const float* gamma, const float* beta, float* running_mean,
                             float* running_var, long long* num_batches_tracked, float momentum, float eps, float* y, float* saved,
                             uint8_t* relu_mask, float* scratch, long long M, int C, int relu, cudaStream_t stream) {
  if (check_c(C)) return -1;
  const int splits = choose_splits(M, C);
  const long long rps = (M + splits - 1) / splits;
  const int cgroups = (C + kCG - 1) / kCG;
  const FinalizeArgs fin{gamma, beta, running_mean, running_var, momentum, eps, saved, num_batches_tracked};
  bn_stats_kernel<<<dim3(cgroups, splits), kThreads, 0, stream>>>(x, scratch, M, C, rps, fin);  // + finalize in each group's last CTA
  apply(x, residual, y, saved + 2ll * C, saved + 3ll * C, M, C, relu, relu_mask, stream);
  PDL_LAUNCH_CHECK();
  return 0;
}

namespace {
// Statistics already reduced to `splits` (count, mean, M2) triples per channel by the producer of x (the tcgen05
// convolution's epilogue): only the per-channel merge is left.
__global__ void __launch_bounds__(kThreads) bn_finalize_only_kernel(const float* __restrict__ partial, int splits, int C, const FinalizeArgs fin) {
  finalize_group(partial, splits, C, blockIdx.x, fin);
}
}  // namespace

// Training forward when the producer of x already left per-channel partial statistics ([splits][C][3]: count, mean,
// M2 -- pdl_tc_conv2d_fwd_bnstats): x is read once (normalise) instead of twice.  Same outputs as pdl_bn_fwd_train.
PDL_API int pdl_bn_fwd_train_prestats(const float* x, const float* residual, const float* gamma, const float* beta, float* running_mean,
                                      float* running_var, long long* num_batches_tracked, float momentum, float eps, float* y, float* saved,
                                      uint8_t* relu_mask, const float* partial, int splits, long long M, int C, int relu, cudaStream_t stream) {
  if (check_c(C)) return -1;
  if (splits < 1) { pdl_set_error("batch norm: no partial statistics"); return -1; }
  const int cgroups = (C + kCG - 1) / kCG;
  const FinalizeArgs fin{gamma, beta, running_mean, running_var, momentum, eps, saved, num_batches_tracked};
  bn_finalize_only_kernel<<<cgroups, kThreads, 0, stream>>>(partial, splits, C, fin);
  apply(x, residual, y, saved + 2ll * C, saved + 3ll * C, M, C, relu, relu_mask, stream);
  PDL_LAUNCH_CHECK();
  return 0;
}

PDL_API int pdl_bn_fwd_eval(const float* x, const float* residual, const float* gamma, const float* beta, const float* running_mean,
                            const float* running_var, float eps, float* y, float* scratch, long long M, int C, int relu, cudaStream_t stream) {
  if (check_c(C)) return -1;
  float* scale = scratch;
  float* shift = scratch + C;
  bn_eval_coeffs_kernel<<<(C + 255) / 256, 256, 0, stream>>>(C, gamma, beta, running_mean, running_var, eps, scale, shift);
  apply(x, residual, y, scale, shift, M, C, relu, nullptr, stream);
  PDL_LAUNCH_CHECK();
  return 0;
}

// Training backward of y = act(bn(x) [+ residual]): dx, dgamma, dbeta and, when dres != nullptr, the gradient
// of the residual input (= dy through the activation).  relu_mask (written by the forward pass) is only read when
// relu != 0 and dres != nullptr; the plain ReLU form recomputes the mask from x.
// dy2 (optional): a second tensor to be ADDED to dy -- the block output feeds both the next block's first convolution and
// its skip connection, and instead of autograd summing the two gradients in a pass of its own (read 2, write 1) this kernel
// reads both where it reads dy anyway.
PDL_API int pdl_bn_bwd2(const float* x, const float* dy, const float* dy2, const uint8_t* relu_mask, const float* gamma, const float* saved,
                        float* dx, float* dres, float* dgamma, float* dbeta, float* scratch, long long M, int C, int relu, cudaStream_t stream) {
  if (check_c(C)) return -1;
  const int splits = choose_splits(M, C);
  const long long rps = (M + splits - 1) / splits;
  const int cgroups = (C + kCG - 1) / kCG;
  if (cgroups > 2048) { pdl_set_error("batch norm: more than 65536 channels"); return -1; }
  float* partial = scratch;
  float* coef = scratch + static_cast<long long>(splits) * C * 3;
  const int mask = !relu ? 0 : (dres ? 2 : 1);
  if (mask == 2 && !relu_mask) { pdl_set_error("batch norm backward: the residual+ReLU form needs the forward pass's mask"); return -1; }
  const dim3 grid(cgroups, splits);
  const BwdFinalizeArgs fin{static_cast<float>(M), gamma, dgamma, dbeta, coef};
  const uint8_t* y = relu_mask;
  const uint8_t* y4 = relu_mask;
  if (mask == 0) bn_bwd_stats_kernel<0><<<grid, kThreads, 0, stream>>>(x, dy, dy2, y, saved, partial, M, C, rps, fin);
  else if (mask == 1) bn_bwd_stats_kernel<1><<<grid, kThreads, 0, stream>>>(x, dy, dy2, y, saved, partial, M, C, rps, fin);
  else bn_bwd_stats_kernel<2><<<grid, kThreads, 0, stream>>>(x, dy, dy2, y, saved, partial, M, C, rps, fin);
  const long long total4 = M * C / 4;
  const unsigned blocks = static_cast<unsigned>(pdl_ceil_div(total4, kThreads));
  const float4 *x4 = reinterpret_cast<const float4*>(x), *g4 = reinterpret_cast<const float4*>(dy), *h4 = reinterpret_cast<const float4*>(dy2);
  float4 *dx4 = reinterpret_cast<float4*>(dx), *dr4 = reinterpret_cast<float4*>(dres);
  if (mask == 0 && !dres) bn_bwd_apply_kernel<0, false><<<blocks, kThreads, 0, stream>>>(x4, g4, h4, y4, dx4, dr4, coef, saved, total4, C / 4);
  else if (mask == 0) bn_bwd_apply_kernel<0, true><<<blocks, kThreads, 0, stream>>>(x4, g4, h4, y4, dx4, dr4, coef, saved, total4, C / 4);
  else if (mask == 1) bn_bwd_apply_kernel<1, false><<<blocks, kThreads, 0, stream>>>(x4, g4, h4, y4, dx4, dr4, coef, saved, total4, C / 4);
  else bn_bwd_apply_kernel<2, true><<<blocks, kThreads, 0, stream>>>(x4, g4, h4, y4, dx4, dr4, coef, saved, total4, C / 4);
  PDL_LAUNCH_CHECK();
  return 0;
}

PDL_API int pdl_bn_bwd(const float* x, const float* dy, const uint8_t* relu_mask, const float* gamma, const float* saved, float* dx,
                       float* dres, float* dgamma, float* dbeta, float* scratch, long long M, int C, int relu, cudaStream_t stream) {
  return pdl_bn_bwd2(x, dy, nullptr, relu_mask, gamma, saved, dx, dres, dgamma, dbeta, scratch, M, C, relu, stream);
}
