// fc1 (2048 -> 10, strict fp32) fused with CrossEntropy, and its backward fused with pool3-backward.
// Reference ops: x.view(-1, 2048) -> fc1 -> CrossEntropyLoss (launch_dist.py:39-40, :58, :87) and
// their autograd.  Activations are NHWC, so the fc weight is used through the NHWC-ordered copy
// `wfp[j][hw*128 + c] = w[j][c*16 + hw]` made by pdl_convnet_pack (coalesced reads, no flatten kernel).
#include "common.cuh"

#include <cfloat>

namespace {

constexpr int kThreads = 256;
constexpr int kClasses = 10;
constexpr int kFeat = 2048;

__device__ unsigned int g_fc_done = 0;
__device__ unsigned int g_fcw_done[8] = {0, 0, 0, 0, 0, 0, 0, 0};

// Two samples per CTA, four warps per sample (512 features each, all 44 float4 loads of a lane in flight at
// once: the kernel is L2-latency-bound at batch 100).  Then per-row loss and dlogits = (softmax - onehot) / N;
// the last CTA reduces the per-row losses in index order into the mean loss.
constexpr int kRowsPerCta = 2;
// One sample of the fc1 backward fused with the backward of pool3 (2x2/2 over 8x8x128) and the ReLU before it:
// da3 = dlogits * Wf on the fly, routed through argmax + ReLU mask into dz3[n][8][8][128]; the CTA's column sums are
// the conv3 bias-gradient partials db3part[n][128].  All kThreads threads of the CTA take part; dl = the sample's
// dlogits in shared memory, cs = kThreads floats of shared scratch.
__device__ __forceinline__ void fc_dgrad_pool3_row(int n, const float* dl, const float* __restrict__ a3, const uint8_t* __restrict__ idx3,
                                                   const float* __restrict__ wfp, float* __restrict__ dz3, float* __restrict__ db3part, float* cs) {
  const int c = threadIdx.x % 128, half = threadIdx.x / 128;  // 2 pooled pixels in flight per pass
  float colsum = 0.f;
#pragma unroll
  for (int pp = half; pp < 16; pp += 2) {
    const int i = pp * 128 + c;
    float g = 0.f;
#pragma unroll
    for (int j = 0; j < kClasses; ++j) g = fmaf(dl[j], wfp[j * kFeat + i], g);
    const long long o = static_cast<long long>(n) * kFeat + i;
    if (!(a3[o] > 0.f)) g = 0.f;  // ReLU mask: the pooled value is a ReLU output
    const int arg = idx3[o], ph = pp / 4, pw = pp % 4;
    colsum += g;
    g = round_tf32(g);  // dz3 is an operand of the tcgen05 dgrad / wgrad
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int h = 2 * ph + q / 2, w = 2 * pw + q % 2;
      dz3[((static_cast<long long>(n) * 8 + h) * 8 + w) * 128 + c] = (q == arg) ? g : 0.f;
    }
  }
  cs[threadIdx.x] = colsum;
  __syncthreads();
  if (threadIdx.x < 128) db3part[n * 128 + threadIdx.x] = cs[threadIdx.x] + cs[threadIdx.x + 128];
  __syncthreads();
}

// fc1 forward + cross-entropy (+ optionally, when dz3 != nullptr, the fc1 dgrad / pool3 backward of the same rows:
// a row's dlogits only depend on that row, so the CTA that produced them can go straight on to dz3).
// kRows samples per CTA (2: forward only, four warps x 512 features per sample; 1: fused form, eight warps x 256
// features, so that the backward part has one CTA per sample like the stand-alone kernel).
template <int kRows>
__global__ void __launch_bounds__(kThreads) fc_ce_kernel(const float* __restrict__ a3, const float* __restrict__ wfp, const float* __restrict__ bf,
                                                         const long long* __restrict__ labels, float* __restrict__ logits,
                                                         float* __restrict__ row_loss, float* __restrict__ loss, float* __restrict__ dlogits, int N,
                                                         const uint8_t* __restrict__ idx3, float* __restrict__ dz3, float* __restrict__ db3part) {
  constexpr int kWpr = (kThreads / 32) / kRows;  // warps per sample
  constexpr int kIters = 16 / kWpr;              // 128-feature float4 sweeps per warp
  __shared__ float part_logit[kRows][kWpr][kClasses];
  __shared__ float sdl[kRows][kClasses + 2];
  __shared__ float cs[kThreads];
  launch_dependents();     // successor may set up while this grid drains
  grid_dependency_wait();  // predecessor's outputs are complete and visible from here on
  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const int row_in_cta = warp / kWpr, quarter = warp % kWpr;
  const int n = blockIdx.x * kRows + row_in_cta;
  float acc[kClasses];
#pragma unroll
  for (int j = 0; j < kClasses; ++j) acc[j] = 0.f;
  if (n < N) {
    const float4* xr = reinterpret_cast<const float4*>(a3 + static_cast<long long>(n) * kFeat) + quarter * (32 * kIters);
    float4 xv[kIters];
#pragma unroll
    for (int it = 0; it < kIters; ++it) xv[it] = xr[it * 32 + lane];
#pragma unroll
    for (int j = 0; j < kClasses; ++j) {
      const float4* wr = reinterpret_cast<const float4*>(wfp + j * kFeat) + quarter * (32 * kIters);
      float4 wv[kIters];
#pragma unroll
      for (int it = 0; it < kIters; ++it) wv[it] = wr[it * 32 + lane];
#pragma unroll
      for (int it = 0; it < kIters; ++it)
        acc[j] = fmaf(xv[it].x, wv[it].x, fmaf(xv[it].y, wv[it].y, fmaf(xv[it].z, wv[it].z, fmaf(xv[it].w, wv[it].w, acc[j]))));
    }
  }
#pragma unroll
  for (int j = 0; j < kClasses; ++j) {
#pragma unroll
    for (int o = 16; o; o >>= 1) acc[j] += __shfl_xor_sync(0xffffffffu, acc[j], o);
  }
  if (lane == 0) {
#pragma unroll
    for (int j = 0; j < kClasses; ++j) part_logit[row_in_cta][quarter][j] = acc[j];
  }
  __syncthreads();
  if (quarter == 0 && n < N) {  // one warp finishes the row; lane j < 10 owns class j
    float mine = -FLT_MAX;
    if (lane < kClasses) {
      float s = part_logit[row_in_cta][0][lane];
#pragma unroll
      for (int q = 1; q < kWpr; ++q) s += part_logit[row_in_cta][q][lane];  // fixed order
      mine = s + bf[lane];
    }
    float mx = mine;
#pragma unroll
    for (int o = 16; o; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
    float ex = lane < kClasses ? expf(mine - mx) : 0.f, sum = ex;
#pragma unroll
    for (int o = 16; o; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
    const float lse = logf(sum) + mx;
    const int tgt = static_cast<int>(labels[n]);
    if (lane < kClasses) {
      logits[n * kClasses + lane] = mine;
      const float d = (expf(mine - lse) - (lane == tgt ? 1.f : 0.f)) / N;
      dlogits[n * kClasses + lane] = d;
      sdl[row_in_cta][lane] = d;
      if (lane == tgt) row_loss[n] = lse - mine;
    }
  }
  if (dz3 != nullptr) {
    __syncthreads();  // sdl is complete
    for (int r = 0; r < kRows; ++r) {
      const int nn = blockIdx.x * kRows + r;
      if (nn >= N) break;  // uniform across the CTA
      fc_dgrad_pool3_row(nn, sdl[r], a3, idx3, wfp, dz3, db3part, cs);
    }
  }
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    is_last = atomicAdd(&g_fc_done, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    __shared__ float part[kThreads];
    float acc = 0.f;
    const int per = (N + kThreads - 1) / kThreads, lo = threadIdx.x * per, hi = min(N, lo + per);
    for (int i = lo; i < hi; ++i) acc += __ldcg(row_loss + i);
    part[threadIdx.x] = acc;
    __syncthreads();
    for (int s = kThreads / 2; s; s >>= 1) {
      if (threadIdx.x < s) part[threadIdx.x] += part[threadIdx.x + s];
      __syncthreads();
    }
    if (threadIdx.x == 0) { *loss = part[0] / N; g_fc_done = 0; }
  }
}

// Backward of fc1 fused with backward of pool3 (2x2/2 over 8x8x128) and the ReLU before it.
//  role A (blockIdx.x <  8*splits): dWf / dbf partial sums over a slice of the batch; the last CTA of each
//         256-column group reduces the slices in order and writes dWf (standard [10][c*16+hw] layout) + dbf.
//  role B (the rest): one CTA per sample: da3 = dlogits * Wf on the fly, routed through argmax + ReLU mask
//         into dz3[n][8][8][128]; per-CTA column sums = conv3 bias-gradient partials db3part[n][128].
__global__ void __launch_bounds__(kThreads) fc_bwd_pool3_kernel(const float* __restrict__ a3, const uint8_t* __restrict__ idx3,
                                                                const float* __restrict__ dlogits, const float* __restrict__ wfp,
                                                                float* __restrict__ gwf, float* __restrict__ gbf, float* __restrict__ wpart,
                                                                float* __restrict__ dz3, float* __restrict__ db3part, int N, int splits,
                                                                int mode) {
  __shared__ float dl[64 * kClasses];
  launch_dependents();     // successor may set up while this grid drains
  grid_dependency_wait();  // predecessor's outputs are complete and visible from here on
  // mode 0: both roles in one grid; 1: role A only (fc weight gradient -- off the critical path, the engine puts it
  // on its branch stream); 2: role B only (dz3 / db3part, which the conv3 dgrad is waiting for)
  const int roleA_blocks = mode == 2 ? 0 : 8 * splits;
  if (blockIdx.x < roleA_blocks) {
    const int colg = blockIdx.x % 8, split = blockIdx.x / 8;
    const int i = colg * kThreads + threadIdx.x;  // NHWC feature index
    const int per = (N + splits - 1) / splits, n_lo = split * per, n_hi = min(N, n_lo + per);
    float acc[kClasses], accb = 0.f;
#pragma unroll
    for (int j = 0; j < kClasses; ++j) acc[j] = 0.f;
    for (int n0 = n_lo; n0 < n_hi; n0 += 64) {
      const int cnt = min(64, n_hi - n0);
      __syncthreads();
      for (int t = threadIdx.x; t < cnt * kClasses; t += kThreads) dl[t] = dlogits[n0 * kClasses + t];
      __syncthreads();
#pragma unroll 8
      for (int m = 0; m < cnt; ++m) {
        const float xv = a3[static_cast<long long>(n0 + m) * kFeat + i];
#pragma unroll
        for (int j = 0; j < kClasses; ++j) acc[j] = fmaf(dl[m * kClasses + j], xv, acc[j]);
        if (colg == 0 && threadIdx.x < kClasses) accb += dl[m * kClasses + threadIdx.x];
      }
    }
#pragma unroll
    for (int j = 0; j < kClasses; ++j) wpart[(static_cast<long long>(split) * kClasses + j) * kFeat + i] = acc[j];
    if (colg == 0 && threadIdx.x < kClasses) wpart[static_cast<long long>(splits) * kClasses * kFeat + split * kClasses + threadIdx.x] = accb;
    __shared__ bool is_last;
    __syncthreads();
    if (threadIdx.x == 0) {
      __threadfence();
      is_last = atomicAdd(&g_fcw_done[colg], 1u) == static_cast<unsigned>(splits) - 1;
    }
    __syncthreads();
    if (is_last) {
      __threadfence();
      const int hw = i / 128, c = i % 128;
#pragma unroll
      for (int j = 0; j < kClasses; ++j) {
        float s = 0.f;
#pragma unroll 8
        for (int sp = 0; sp < splits; ++sp) s += __ldcg(wpart + (static_cast<long long>(sp) * kClasses + j) * kFeat + i);
        gwf[j * kFeat + c * 16 + hw] = s;
      }
      if (colg == 0 && threadIdx.x < kClasses) {
        float s = 0.f;
        for (int sp = 0; sp < splits; ++sp) s += __ldcg(wpart + static_cast<long long>(splits) * kClasses * kFeat + sp * kClasses + threadIdx.x);
        gbf[threadIdx.x] = s;
      }
      if (threadIdx.x == 0) g_fcw_done[colg] = 0;
    }
    return;
  }
  // ---- role B ----
  const int n = blockIdx.x - roleA_blocks;
  if (n >= N || mode == 1) return;
  if (threadIdx.x < kClasses) dl[threadIdx.x] = dlogits[n * kClasses + threadIdx.x];
  __syncthreads();
  __shared__ float cs[kThreads];
  fc_dgrad_pool3_row(n, dl, a3, idx3, wfp, dz3, db3part, cs);
}

}  // namespace

PDL_API int pdl_fc_ce_fused(const float* a3, const float* wfp, const float* bf, const long long* labels, float* logits, float* row_loss, float* loss,
                            float* dlogits, int N, cudaStream_t stream) {
  PDL_CUDA_CHECK(launch_dependent(fc_ce_kernel<kRowsPerCta>, dim3(pdl_ceil_div(N, kRowsPerCta)), dim3(kThreads), 0, stream, a3, wfp, bf, labels, logits,
                                  row_loss, loss, dlogits, N, static_cast<const uint8_t*>(nullptr), static_cast<float*>(nullptr),
                                  static_cast<float*>(nullptr)));
  PDL_LAUNCH_CHECK();
  return 0;
}

// fc1 forward + loss + dlogits AND, in the same kernel, part 2 of pdl_fc_bwd_pool3_fused (dz3, db3part): the
// critical path of the step loses one kernel boundary.  The fc1 weight gradient (part 1) still needs every row's
// dlogits and stays a separate launch.
PDL_API int pdl_fc_ce_bwd_fused(const float* a3, const float* wfp, const float* bf, const long long* labels, float* logits, float* row_loss,
                                float* loss, float* dlogits, const uint8_t* idx3, float* dz3, float* db3part, int N, cudaStream_t stream) {
  PDL_CUDA_CHECK(launch_dependent(fc_ce_kernel<1>, dim3(N), dim3(kThreads), 0, stream, a3, wfp, bf, labels, logits, row_loss, loss, dlogits, N, idx3,
                                  dz3, db3part));
  PDL_LAUNCH_CHECK();
  return 0;
}

PDL_API int pdl_fc_bwd_splits(int N) { int s = (N + 15) / 16; return s < 1 ? 1 : (s > 32 ? 32 : s); }
// floats of scratch pdl_fc_bwd_pool3_fused needs
PDL_API long long pdl_fc_bwd_scratch_floats(int N) { long long s = pdl_fc_bwd_splits(N); return s * kClasses * kFeat + s * kClasses; }

PDL_API int pdl_fc_bwd_pool3_fused(const float* a3, const uint8_t* idx3, const float* dlogits, const float* wfp, float* gwf, float* gbf,
                                   float* wpart, float* dz3, float* db3part, int N, cudaStream_t stream) {
  const int splits = pdl_fc_bwd_splits(N);
  PDL_CUDA_CHECK(launch_dependent(fc_bwd_pool3_kernel, dim3(8 * splits + N), dim3(kThreads), 0, stream, a3, idx3, dlogits, wfp, gwf, gbf, wpart, dz3,
                                  db3part, N, splits, 0));
  PDL_LAUNCH_CHECK();
  return 0;
}

// The two halves of pdl_fc_bwd_pool3_fused as separate launches: part 1 = fc1 weight/bias gradient (gwf, gbf),
// part 2 = dz3 + conv3 bias-gradient partials.  Lets a caller overlap part 1 with the rest of the backward pass.
PDL_API int pdl_fc_bwd_pool3_part(int part, const float* a3, const uint8_t* idx3, const float* dlogits, const float* wfp, float* gwf, float* gbf,
                                  float* wpart, float* dz3, float* db3part, int N, cudaStream_t stream) {
  if (part != 1 && part != 2) { pdl_set_error("fc_bwd_pool3_part: part must be 1 or 2"); return -1; }
  const int splits = pdl_fc_bwd_splits(N);
  const int blocks = part == 1 ? 8 * splits : N;
  PDL_CUDA_CHECK(launch_dependent(fc_bwd_pool3_kernel, dim3(blocks), dim3(kThreads), 0, stream, a3, idx3, dlogits, wfp, gwf, gbf, wpart, dz3, db3part, N,
                                  splits, part));
  PDL_LAUNCH_CHECK();
  return 0;
}
