// BatchNorm2d for the ResNets the reference trains (torchvision resnet18 at example_launch.py:26; ResNet-50
// per BASELINE.json), NHWC fp32.  Training mode: per-channel batch statistics (single pass, Welford, merged
// in fixed order => deterministic), normalise + affine (+ optional fused ReLU), running-stat update with the
// unbiased variance exactly like torch.  Backward: dgamma/dbeta and dx in the standard closed form.
//   x, y, dy, dx : [M][C] with M = N*H*W rows, channels contiguous
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kCG = 32;              // channels per CTA (one 128-byte row segment)
constexpr int kRL = kThreads / kCG;  // row lanes per CTA

struct Welford { float n, mean, m2; };
__device__ __forceinline__ Welford merge(Welford a, Welford b) {
  if (b.n == 0.f) return a;
  if (a.n == 0.f) return b;
  const float n = a.n + b.n, d = b.mean - a.mean;
  return {n, a.mean + d * (b.n / n), a.m2 + b.m2 + d * d * (a.n * b.n / n)};
}

// partial[(split * C + c) * 3 + {0,1,2}] = (count, mean, M2) of x[rows of this split][c]
__global__ void __launch_bounds__(kThreads) bn_stats_kernel(const float* __restrict__ x, float* __restrict__ partial, long long M, int C,
                                                            long long rows_per_split) {
  const int cl = threadIdx.x % kCG, rl = threadIdx.x / kCG;
  const int c = blockIdx.x * kCG + cl;
  const long long r_lo = blockIdx.y * rows_per_split, r_hi = min(M, r_lo + rows_per_split);
  Welford w{0.f, 0.f, 0.f};
  if (c < C) {
#pragma unroll 4
    for (long long r = r_lo + rl; r < r_hi; r += kRL) {
      const float v = x[r * C + c];
      w.n += 1.f;
      const float d = v - w.mean;
      w.mean += d / w.n;
      w.m2 += d * (v - w.mean);
    }
  }
  __shared__ Welford sh[kRL][kCG];
  sh[rl][cl] = w;
  __syncthreads();
  if (rl == 0 && c < C) {
    Welford t = sh[0][cl];
#pragma unroll
    for (int i = 1; i < kRL; ++i) t = merge(t, sh[i][cl]);
    float* o = partial + (static_cast<long long>(blockIdx.y) * C + c) * 3;
    o[0] = t.n; o[1] = t.mean; o[2] = t.m2;
  }
}

// Merge the splits per channel; emit mean / invstd / scale / shift; update running stats (torch semantics).
__global__ void bn_finalize_kernel(const float* __restrict__ partial, int splits, int C, const float* __restrict__ gamma,
                                   const float* __restrict__ beta, float* __restrict__ running_mean, float* __restrict__ running_var,
                                   float momentum, float eps, float* __restrict__ save_mean, float* __restrict__ save_invstd,
                                   float* __restrict__ scale, float* __restrict__ shift) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  Welford t{0.f, 0.f, 0.f};
  for (int s = 0; s < splits; ++s) {
    const float* p = partial + (static_cast<long long>(s) * C + c) * 3;
    t = merge(t, Welford{p[0], p[1], p[2]});
  }
  const float var = t.m2 / t.n;  // biased: used for normalisation
  const float invstd = rsqrtf(var + eps);
  save_mean[c] = t.mean;
  save_invstd[c] = invstd;
  const float g = gamma ? gamma[c] : 1.f, b = beta ? beta[c] : 0.f;
  scale[c] = g * invstd;
  shift[c] = b - t.mean * g * invstd;
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
  shift[c] = b - rm[c] * g * invstd;
}

// y = x * scale[c] + shift[c]  (+ ReLU); C % 4 == 0
__global__ void __launch_bounds__(kThreads) bn_apply_kernel(const float4* __restrict__ x, float4* __restrict__ y, const float* __restrict__ scale,
                                                            const float* __restrict__ shift, long long total4, int C4, int relu) {
  const long long i = blockIdx.x * static_cast<long long>(kThreads) + threadIdx.x;
  if (i >= total4) return;
  const int c = static_cast<int>(i % C4) * 4;
  const float4 v = x[i], s = *reinterpret_cast<const float4*>(scale + c), h = *reinterpret_cast<const float4*>(shift + c);
  float4 o = make_float4(fmaf(v.x, s.x, h.x), fmaf(v.y, s.y, h.y), fmaf(v.z, s.z, h.z), fmaf(v.w, s.w, h.w));
  if (relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
  y[i] = o;
}

// partial[(split * C + c) * 2 + {0,1}] = (sum dy, sum dy * xhat) over the rows of the split
__global__ void __launch_bounds__(kThreads) bn_bwd_stats_kernel(const float* __restrict__ x, const float* __restrict__ dy,
                                                                const float* __restrict__ mean, const float* __restrict__ invstd,
                                                                float* __restrict__ partial, long long M, int C, long long rows_per_split) {
  const int cl = threadIdx.x % kCG, rl = threadIdx.x / kCG;
  const int c = blockIdx.x * kCG + cl;
  const long long r_lo = blockIdx.y * rows_per_split, r_hi = min(M, r_lo + rows_per_split);
  float s1 = 0.f, s2 = 0.f;
  if (c < C) {
    const float mu = mean[c], is = invstd[c];
#pragma unroll 4
    for (long long r = r_lo + rl; r < r_hi; r += kRL) {
      const float g = dy[r * C + c];
      s1 += g;
      s2 = fmaf(g, (x[r * C + c] - mu) * is, s2);
    }
  }
  __shared__ float a1[kRL][kCG], a2[kRL][kCG];
  a1[rl][cl] = s1; a2[rl][cl] = s2;
  __syncthreads();
  if (rl == 0 && c < C) {
    float t1 = 0.f, t2 = 0.f;
#pragma unroll
    for (int i = 0; i < kRL; ++i) { t1 += a1[i][cl]; t2 += a2[i][cl]; }
    float* o = partial + (static_cast<long long>(blockIdx.y) * C + c) * 2;
    o[0] = t1; o[1] = t2;
  }
}

// dgamma, dbeta and the per-channel coefficients of dx = a*dy + b*x + d
__global__ void bn_bwd_finalize_kernel(const float* __restrict__ partial, int splits, int C, float M, const float* __restrict__ gamma,
                                       const float* __restrict__ mean, const float* __restrict__ invstd, float* __restrict__ dgamma,
                                       float* __restrict__ dbeta, float* __restrict__ ca, float* __restrict__ cb, float* __restrict__ cd) {
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= C) return;
  float s1 = 0.f, s2 = 0.f;
  for (int s = 0; s < splits; ++s) { s1 += partial[(static_cast<long long>(s) * C + c) * 2]; s2 += partial[(static_cast<long long>(s) * C + c) * 2 + 1]; }
  if (dgamma) dgamma[c] = s2;
  if (dbeta) dbeta[c] = s1;
  // dx = g*is * (dy - s1/M - xhat * s2/M),  xhat = (x - mu) * is
  const float g = gamma ? gamma[c] : 1.f, is = invstd[c], mu = mean[c];
  const float a = g * is, k = a * is * s2 / M;
  ca[c] = a;
  cb[c] = -k;
  cd[c] = -a * s1 / M + k * mu;
}

__global__ void __launch_bounds__(kThreads) bn_bwd_apply_kernel(const float4* __restrict__ x, const float4* __restrict__ dy, float4* __restrict__ dx,
                                                                const float* __restrict__ ca, const float* __restrict__ cb,
                                                                const float* __restrict__ cd, long long total4, int C4) {
  const long long i = blockIdx.x * static_cast<long long>(kThreads) + threadIdx.x;
  if (i >= total4) return;
  const int c = static_cast<int>(i % C4) * 4;
  const float4 v = x[i], g = dy[i];
  const float4 a = *reinterpret_cast<const float4*>(ca + c), b = *reinterpret_cast<const float4*>(cb + c), d = *reinterpret_cast<const float4*>(cd + c);
  dx[i] = make_float4(fmaf(a.x, g.x, fmaf(b.x, v.x, d.x)), fmaf(a.y, g.y, fmaf(b.y, v.y, d.y)), fmaf(a.z, g.z, fmaf(b.z, v.z, d.z)),
                      fmaf(a.w, g.w, fmaf(b.w, v.w, d.w)));
}

int choose_splits(long long M, int C) {
  const int cgroups = (C + kCG - 1) / kCG;
  long long s = (4 * 148 + cgroups - 1) / cgroups;
  const long long by_rows = (M + 63) / 64;
  if (s > by_rows) s = by_rows;
  if (s < 1) s = 1;
  if (s > 64) s = 64;  // the finalize kernel merges the splits serially per channel
  return static_cast<int>(s);
}

}  // namespace

// floats of scratch the BN entry points need: split partials + 5 per-channel coefficient rows
PDL_API long long pdl_bn_scratch_floats(long long M, int C) { return static_cast<long long>(choose_splits(M, C)) * C * 3 + 5ll * C; }

// Training forward.  save_mean / save_invstd [C] are kept for backward.  relu != 0 fuses the activation.
PDL_API int pdl_bn_fwd_train(const float* x, const float* gamma, const float* beta, float* running_mean, float* running_var, float momentum,
                             float eps, float* y, float* save_mean, float* save_invstd, float* scratch, long long M, int C, int relu,
                             cudaStream_t stream) {
  if (C % 4) { pdl_set_error("batch norm: channel count %d is not a multiple of 4", C); return -1; }
  const int splits = choose_splits(M, C);
  const long long rps = (M + splits - 1) / splits;
  float* partial = scratch;
  float* scale = scratch + static_cast<long long>(splits) * C * 3;
  float* shift = scale + C;
  bn_stats_kernel<<<dim3((C + kCG - 1) / kCG, splits), kThreads, 0, stream>>>(x, partial, M, C, rps);
  bn_finalize_kernel<<<(C + 255) / 256, 256, 0, stream>>>(partial, splits, C, gamma, beta, running_mean, running_var, momentum, eps, save_mean,
                                                          save_invstd, scale, shift);
  const long long total4 = M * C / 4;
  bn_apply_kernel<<<pdl_ceil_div(total4, kThreads), kThreads, 0, stream>>>(reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(y), scale,
                                                                           shift, total4, C / 4, relu);
  PDL_LAUNCH_CHECK();
  return 0;
}

PDL_API int pdl_bn_fwd_eval(const float* x, const float* gamma, const float* beta, const float* running_mean, const float* running_var, float eps,
                            float* y, float* scratch, long long M, int C, int relu, cudaStream_t stream) {
  if (C % 4) { pdl_set_error("batch norm: channel count %d is not a multiple of 4", C); return -1; }
  float* scale = scratch;
  float* shift = scratch + C;
  bn_eval_coeffs_kernel<<<(C + 255) / 256, 256, 0, stream>>>(C, gamma, beta, running_mean, running_var, eps, scale, shift);
  const long long total4 = M * C / 4;
  bn_apply_kernel<<<pdl_ceil_div(total4, kThreads), kThreads, 0, stream>>>(reinterpret_cast<const float4*>(x), reinterpret_cast<float4*>(y), scale,
                                                                           shift, total4, C / 4, relu);
  PDL_LAUNCH_CHECK();
  return 0;
}

// Training backward: dx, dgamma, dbeta from x, dy and the saved statistics.
PDL_API int pdl_bn_bwd(const float* x, const float* dy, const float* gamma, const float* save_mean, const float* save_invstd, float* dx,
                       float* dgamma, float* dbeta, float* scratch, long long M, int C, cudaStream_t stream) {
  if (C % 4) { pdl_set_error("batch norm: channel count %d is not a multiple of 4", C); return -1; }
  const int splits = choose_splits(M, C);
  const long long rps = (M + splits - 1) / splits;
  float* partial = scratch;
  float* ca = scratch + static_cast<long long>(splits) * C * 3;
  float* cb = ca + C;
  float* cd = cb + C;
  bn_bwd_stats_kernel<<<dim3((C + kCG - 1) / kCG, splits), kThreads, 0, stream>>>(x, dy, save_mean, save_invstd, partial, M, C, rps);
  bn_bwd_finalize_kernel<<<(C + 255) / 256, 256, 0, stream>>>(partial, splits, C, static_cast<float>(M), gamma, save_mean, save_invstd, dgamma, dbeta,
                                                              ca, cb, cd);
  const long long total4 = M * C / 4;
  bn_bwd_apply_kernel<<<pdl_ceil_div(total4, kThreads), kThreads, 0, stream>>>(reinterpret_cast<const float4*>(x), reinterpret_cast<const float4*>(dy),
                                                                               reinterpret_cast<float4*>(dx), ca, cb, cd, total4, C / 4);
  PDL_LAUNCH_CHECK();
  return 0;
}
