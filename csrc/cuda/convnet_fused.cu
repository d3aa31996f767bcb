// Fused CUDA-core kernels for the parts of the reference's ConvNet (launch_dist.py:9-41) that are
// not GEMM-shaped enough for the tensor cores, written so the whole training step is ~16 launches:
//   * weight repack for the tcgen05 conv kernels + NHWC-ordered copy of the fc weight   (1 launch)
//   * conv1 (Cin = 1, K = 25) + bias + ReLU + 2x2/2 max-pool with argmax                 (1 launch)
//   * fc1 + CrossEntropy forward AND d(loss)/d(logits)                                   (1 launch)
//   * fc1 wgrad/bias-grad + fc1 dgrad folded into pool3-backward (+ conv3 bias partials) (1 launch)
//   * overlapping-window pool backward with ReLU mask + bias-gradient partials           (1 launch)
//   * conv1 wgrad with pool1-backward + ReLU mask folded into the operand load           (1 launch)
// All reductions are fixed-order (deterministic); multi-CTA reductions finish in the last CTA.
#include "common.cuh"

#include <cfloat>

namespace {

constexpr int kThreads = 256;

// ------------------------------------------------------------------------------------------------
// 1. weight repack: w[co][ci][tap] -> fwd[tap][co][ci], bwd[tap][ci][co]; fc w[j][c*16+hw] -> [j][hw*128+c]
// ------------------------------------------------------------------------------------------------
struct PackArgs {
  const float *w2, *w3, *wf;
  float *p2f, *p2b, *p3f, *p3b, *wfp;
};

__global__ void __launch_bounds__(kThreads) convnet_pack_kernel(PackArgs a) {
  const int n2 = 64 * 32 * 9, n3 = 128 * 64 * 9, nf = 10 * 2048;
  int i = blockIdx.x * kThreads + threadIdx.x;
  if (i < n2) {
    int tap = i % 9, ci = (i / 9) % 32, co = i / (9 * 32);
    float v = round_tf32(a.w2[i]);
    a.p2f[(tap * 64 + co) * 32 + ci] = v;
    a.p2b[(tap * 32 + ci) * 64 + co] = v;
    return;
  }
  i -= n2;
  if (i < n3) {
    int tap = i % 9, ci = (i / 9) % 64, co = i / (9 * 64);
    float v = round_tf32(a.w3[i]);
    a.p3f[(tap * 128 + co) * 64 + ci] = v;
    a.p3b[(tap * 64 + ci) * 128 + co] = v;
    return;
  }
  i -= n3;
  if (i < nf) {
    int j = i / 2048, k = i % 2048, c = k / 16, hw = k % 16;
    a.wfp[j * 2048 + hw * 128 + c] = a.wf[i];
  }
}

// ------------------------------------------------------------------------------------------------
// 2. conv1: x[N][28][28] (Cin = 1), w[32][25], pad 1 -> conv 26x26x32 -> ReLU -> 2x2/2 pool -> a1[N][13][13][32]
//    One CTA per image: the padded image and the filters sit in shared memory; thread = (channel, pooled
//    pixel), 25 filter taps in registers, the 6x6 input patch is a warp-wide broadcast.
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(kThreads) conv1_fused_kernel(const float* __restrict__ x, const float* __restrict__ w, const float* __restrict__ bias,
                                                               float* __restrict__ a1, uint8_t* __restrict__ idx1, int N) {
  __shared__ float img[30][32];  // 28 + 2 halo, row pitch 32
  launch_dependents();  // conv2's CTAs may set up (barriers, TMEM) while this grid drains; they wait before reading a1
  const int n = blockIdx.x;
  if (n >= N) return;
  for (int i = threadIdx.x; i < 30 * 32; i += kThreads) {
    int r = i / 32, c = i % 32;
    int h = r - 1, ww = c - 1;
    img[r][c] = (h >= 0 && h < 28 && ww >= 0 && ww < 28) ? x[(static_cast<long long>(n) * 28 + h) * 28 + ww] : 0.f;
  }
  const int c = threadIdx.x % 32, pg = threadIdx.x / 32;
  float wr[25];
#pragma unroll
  for (int t = 0; t < 25; ++t) wr[t] = w[c * 25 + t];
  const float b = bias[c];
  __syncthreads();
  for (int pp = pg; pp < 169; pp += kThreads / 32) {
    const int ph = pp / 13, pw = pp % 13;
    float patch[6][6];
#pragma unroll
    for (int i = 0; i < 6; ++i)
#pragma unroll
      for (int j = 0; j < 6; ++j) patch[i][j] = img[2 * ph + i][2 * pw + j];
    float best = 0.f;
    int arg = 0;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      const int dy = q / 2, dx = q % 2;
      float acc = b;
#pragma unroll
      for (int r = 0; r < 5; ++r)
#pragma unroll
        for (int s = 0; s < 5; ++s) acc = fmaf(patch[dy + r][dx + s], wr[r * 5 + s], acc);
      acc = fmaxf(acc, 0.f);
      if (q == 0 || acc > best) { best = acc; arg = q; }
    }
    const long long o = (static_cast<long long>(n) * 169 + pp) * 32 + c;
    a1[o] = round_tf32(best);  // operand of the tcgen05 conv2 forward / wgrad
    idx1[o] = static_cast<uint8_t>(arg);
  }
}

}  // namespace

PDL_API int pdl_convnet_pack(const float* w2, const float* w3, const float* wf, float* p2f, float* p2b, float* p3f, float* p3b, float* wfp,
                             cudaStream_t stream) {
  PackArgs a{w2, w3, wf, p2f, p2b, p3f, p3b, wfp};
  const int total = 64 * 32 * 9 + 128 * 64 * 9 + 10 * 2048;
  convnet_pack_kernel<<<pdl_ceil_div(total, kThreads), kThreads, 0, stream>>>(a);
  PDL_LAUNCH_CHECK();
  return 0;
}

PDL_API int pdl_conv1_fused_fwd(const float* x, const float* w, const float* bias, float* a1, uint8_t* idx1, int N, cudaStream_t stream) {
  if (N > 0) conv1_fused_kernel<<<N, kThreads, 0, stream>>>(x, w, bias, a1, idx1, N);
  PDL_LAUNCH_CHECK();
  return 0;
}
