// Backward-side fused kernels of the ConvNet step (reference autograd of launch_dist.py:36-38):
//   * 2x2 max-pool backward (stride 1 = overlapping windows, or 2) with the preceding ReLU folded in
//     and the bias gradient of the preceding conv produced as per-CTA partial sums;
//   * conv1 weight/bias gradient with pool1-backward + ReLU folded into the operand load
//     (Cin = 1, 25 taps: CUDA cores; dz1 [N][26][26][32] is never materialised);
//   * the same kernel finishing with the cross-GPU exchange of that gradient and the SGD update (the reference's
//     last gradient -> bucket allreduce -> optimizer.step(), launch_dist.py:91-92): conv1's gradient is the LAST one
//     backward produces, so whatever reduces it is fully exposed; here the CTA that holds the finished 832 values
//     pushes them into every peer's inbox over NVLink, raises one flag, and sums the W inbox rows it received.
// Both are latency-bound at batch 100, so loads are issued branch-free / staged through shared
// memory to keep many of them in flight.
#include "comm.cuh"

namespace {

constexpr int kThreads = 256;

// dz[n][h][w][c] = sum over the <=4 windows containing (h,w) whose argmax is (h,w) and whose pooled value > 0.
// A CTA owns `pix_per_cta` consecutive pixels of the flattened [N*H*W] axis; column sums -> dbpart[cta][C].
template <int C, int stride>
__global__ void __launch_bounds__(kThreads) pool2_bwd_dbias_kernel(const float* __restrict__ da, const uint8_t* __restrict__ idx,
                                                                   const float* __restrict__ pooled, float* __restrict__ dz,
                                                                   float* __restrict__ dbpart, unsigned total_pix, unsigned H, unsigned W,
                                                                   unsigned pix_per_cta) {
  // A thread owns FOUR channels of a pixel (float4 / uchar4 accesses): C/4 threads per pixel, 1024/C pixels in flight per
  // CTA, so a CTA's pixel range takes 4x fewer serial iterations than with one channel per thread -- this kernel sits on the
  // step's critical path and is pure latency (13.5 us -> see profiles/convnet_timeline_w1.txt).
  const int PH = (static_cast<int>(H) - 2) / stride + 1, PW = (static_cast<int>(W) - 2) / stride + 1;
  constexpr int C4 = C / 4;            // threads per pixel
  constexpr int G = kThreads / C4;     // pixel lanes
  const int q = threadIdx.x % C4, grp = threadIdx.x / C4;
  float4 colsum = make_float4(0.f, 0.f, 0.f, 0.f);
  launch_dependents();     // successor may set up while this grid drains
  grid_dependency_wait();  // predecessor's outputs are complete and visible from here on
  const unsigned p_lo = blockIdx.x * pix_per_cta;  // 32-bit index math: 64-bit div/mod made this kernel ALU-bound
  const unsigned p_hi = min(total_pix, p_lo + pix_per_cta);
  const float4* da4 = reinterpret_cast<const float4*>(da);
  const float4* po4 = reinterpret_cast<const float4*>(pooled);
  const uchar4* ix4 = reinterpret_cast<const uchar4*>(idx);
#pragma unroll 2
  for (unsigned px = p_lo + grp; px < p_hi; px += G) {
    const int w = static_cast<int>(px % W);
    const unsigned t = px / W;
    const int h = static_cast<int>(t % H);
    const int n = static_cast<int>(t / H);
    float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
    for (int k = 0; k < 4; ++k) {  // window whose element (r,s) = (k/2, k%2) is this pixel
      const int r = k >> 1, s2 = k & 1;
      const int ph_ = h - r, pw_ = w - s2;
      const bool valid = ph_ >= 0 && pw_ >= 0 && (ph_ % stride) == 0 && (pw_ % stride) == 0 && (ph_ / stride) < PH && (pw_ / stride) < PW;
      const int o = valid ? ((n * PH + ph_ / stride) * PW + pw_ / stride) * C4 + q : 0;  // branch-free: loads overlap
      const float4 g = da4[o], pv = po4[o];
      const uchar4 a = ix4[o];
      acc.x += (valid && a.x == k && pv.x > 0.f) ? g.x : 0.f;
      acc.y += (valid && a.y == k && pv.y > 0.f) ? g.y : 0.f;
      acc.z += (valid && a.z == k && pv.z > 0.f) ? g.z : 0.f;
      acc.w += (valid && a.w == k && pv.w > 0.f) ? g.w : 0.f;
    }
    colsum.x += acc.x; colsum.y += acc.y; colsum.z += acc.z; colsum.w += acc.w;
    // operand of the tcgen05 dgrad / wgrad: rounded to TF32 by its producer
    reinterpret_cast<float4*>(dz)[static_cast<size_t>(px) * C4 + q] = make_float4(round_tf32(acc.x), round_tf32(acc.y), round_tf32(acc.z), round_tf32(acc.w));
  }
  __shared__ float4 cs[kThreads];
  cs[threadIdx.x] = colsum;
  __syncthreads();
  if (threadIdx.x < C) {  // channel c = 4 * q + j: sum the G pixel lanes in lane order (deterministic)
    const int qq = threadIdx.x / 4, j = threadIdx.x % 4;
    float s = 0.f;
#pragma unroll
    for (int g = 0; g < G; ++g) s += reinterpret_cast<const float*>(&cs[g * C4 + qq])[j];
    dbpart[blockIdx.x * C + threadIdx.x] = s;
  }
}

__device__ unsigned int g_c1w_done = 0;

// conv1 wgrad: gw1[c][r][s] = sum_{n,ph,pw} g * x[n][2ph+dy+r-1][2pw+dx+s-1],  g = (a1 > 0) ? da1 : 0,
// (dy,dx) = argmax of the pool window.  gb1[c] = sum g.  One CTA (16 warps) per image at a time: image, g and
// argmax are staged in shared memory with coalesced independent loads; lane = channel, warp = slice of the
// 169 pooled pixels, all 25 taps accumulate in registers (index math amortised over 25 FMAs).  The 16 warps
// are summed in shared memory in warp order, the CTA's row goes to `partial`, the last CTA adds the rows in
// CTA order (deterministic).
constexpr int kC1Threads = 512, kC1Warps = kC1Threads / 32;
// Image rows are 34 floats apart: the four arg-max offsets {0, 1, 34, 35} a warp's lanes can ask for then sit in four
// different banks (with a stride of 32, rows r and r+1 collide and every tap load is a 2-way conflict).
constexpr int kC1ImgLd = 34;
constexpr int kC1StageFloats = 30 * kC1ImgLd + 169 * 32 + 169 * 32 / 4;  // img + g + argmax bytes

// Cross-GPU tail of the kernel (kComm): a push-style one-shot allreduce of 832 floats + the SGD update.
//   inbox[q]  rank q's inbox in symmetric memory: [2 parities][W source ranks][832]; a rank writes its row into EVERY
//             peer's inbox (fire-and-forget stores), then raises flag[rank] in every peer's pad with st.release.sys.
//   The receiver polls its LOCAL pad, then reads its LOCAL inbox: no remote load is ever waited for, so the exposed
//   cost is one NVLink store latency plus the skew between ranks.
//   One flag round per step is enough because rows alternate between two parities: a rank overwrites parity e&1 again
//   only in step e+2, after it has seen every peer's flag of step e+1 -- which a peer raises from a kernel that is
//   stream-ordered after its step-e reads.
struct TailComm {
  float* inbox[pdl_comm::kMaxWorld];
  uint32_t* pad[pdl_comm::kMaxWorld];
  pdl_comm::CommState* state;
  float* avg_w;   // averaged gradient out (what param.grad shows), 800 + 32
  float* avg_b;
  float* param_w; // parameters / momentum (nullptr: no update)
  float* param_b;
  float* mom_w;
  float* mom_b;
  SgdHyper* hyper;
  int rank, world;
  float scale;
  unsigned long long timeout_ns;
};

template <bool kComm>
__global__ void __launch_bounds__(kC1Threads) conv1_wgrad_fused_kernel(const float* __restrict__ x, const float* __restrict__ da1,
                                                                       const uint8_t* __restrict__ idx1, const float* __restrict__ a1,
                                                                       float* __restrict__ gw1, float* __restrict__ gb1,
                                                                       float* __restrict__ partial, int N, int imgs_per_cta, TailComm tc) {
  extern __shared__ float c1_smem[];
  float* img = c1_smem;               // [30][kC1ImgLd]
  float* sg = img + 30 * kC1ImgLd;    // [169][32]
  uint8_t* sarg = reinterpret_cast<uint8_t*>(sg + 169 * 32);
  float* red = c1_smem + kC1StageFloats;  // [16][832]
  const int c = threadIdx.x % 32, wp = threadIdx.x / 32;
  float acc[25], accb = 0.f;
#pragma unroll
  for (int t = 0; t < 25; ++t) acc[t] = 0.f;
  launch_dependents();     // successor may set up while this grid drains
  grid_dependency_wait();  // predecessor's outputs are complete and visible from here on
  const int n_lo = blockIdx.x * imgs_per_cta, n_hi = min(N, n_lo + imgs_per_cta);
  for (int n = n_lo; n < n_hi; ++n) {
    __syncthreads();
    for (int i = threadIdx.x; i < 30 * 32; i += kC1Threads) {
      const int r = i / 32, cc = i % 32, h = r - 1, w = cc - 1;
      img[r * kC1ImgLd + cc] = (h >= 0 && h < 28 && w >= 0 && w < 28) ? x[(static_cast<long long>(n) * 28 + h) * 28 + w] : 0.f;
    }
    const long long base = static_cast<long long>(n) * 169 * 32;
#pragma unroll 11
    for (int i = threadIdx.x; i < 169 * 32; i += kC1Threads) {
      const float a = a1[base + i], d = da1[base + i];
      sg[i] = a > 0.f ? d : 0.f;
      sarg[i] = idx1[base + i];
    }
    __syncthreads();
    for (int pp = wp; pp < 169; pp += kC1Warps) {
      const float g = sg[pp * 32 + c];
      const int arg = sarg[pp * 32 + c];
      const float* patch = img + (2 * (pp / 13) + (arg >> 1)) * kC1ImgLd + 2 * (pp % 13) + (arg & 1);
      accb += g;
#pragma unroll
      for (int t = 0; t < 25; ++t) acc[t] = fmaf(g, patch[(t / 5) * kC1ImgLd + t % 5], acc[t]);
    }
  }
#pragma unroll
  for (int t = 0; t < 25; ++t) red[wp * 832 + c * 25 + t] = acc[t];
  red[wp * 832 + 800 + c] = accb;
  __syncthreads();
  for (int i = threadIdx.x; i < 832; i += kC1Threads) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < kC1Warps; ++w) s += red[w * 832 + i];
    partial[static_cast<size_t>(blockIdx.x) * 832 + i] = s;
  }
  __shared__ bool is_last;
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    is_last = atomicAdd(&g_c1w_done, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (is_last) {
    __threadfence();
    // 208 float4 columns x 2 halves of the row list = 416 threads, 16 independent loads in flight each; the two halves
    // are then added in a fixed order (deterministic)
    const unsigned nb = gridDim.x, half_rows = (nb + 1) / 2;
    float4* red4 = reinterpret_cast<float4*>(red);
    if (threadIdx.x < 416) {
      const unsigned col4 = threadIdx.x % 208, half = threadIdx.x / 208;
      const unsigned b_lo = half * half_rows, b_hi = min(nb, b_lo + half_rows);
      float4 s = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 16
      for (unsigned b = b_lo; b < b_hi; ++b) {
        const float4 v = __ldcg(reinterpret_cast<const float4*>(partial + static_cast<size_t>(b) * 832) + col4);
        s.x += v.x; s.y += v.y; s.z += v.z; s.w += v.w;
      }
      red4[half * 208 + col4] = s;
    }
    __syncthreads();
    if (threadIdx.x == 0) g_c1w_done = 0;
    if (!kComm) {
      for (int i = threadIdx.x; i < 832; i += kC1Threads) {
        const float s = red[i] + red[832 + i];
        if (i < 800) gw1[i] = s; else gb1[i - 800] = s;
      }
      return;
    }
    using namespace pdl_comm;
    const int W = tc.world, r = tc.rank;
    const uint32_t e = W > 1 ? tc.state->epoch[0] : 0u;  // no communicator state without peers
    // sticky: an earlier step lost a peer -> leave everything as it is (decided once for the whole CTA)
    const bool dead = __syncthreads_or(W > 1 && tc.state->error != 0) != 0;
    float* row = red + 2 * 832;              // this rank's finished gradient (also the W == 1 "sum")
    for (int i = threadIdx.x; i < 832; i += kC1Threads) {
      const float s = red[i] + red[832 + i];
      row[i] = s;
      if (i < 800) gw1[i] = s; else gb1[i - 800] = s;  // raw local gradient (bucket `in` view)
    }
    __syncthreads();
    if (dead) return;
    float4 sum = make_float4(0.f, 0.f, 0.f, 0.f);
    if (W > 1) {
      const size_t slot = (static_cast<size_t>(e & 1u) * W + r) * 832;
      const float4* row4 = reinterpret_cast<const float4*>(row);
      for (int i = threadIdx.x; i < 208 * W; i += kC1Threads) {  // 208 float4 to each of the W inboxes (own included)
        const int q = i / 208, j = i % 208;
        st_f4(tc.inbox[q] + slot, j, row4[j]);
      }
      __syncthreads();
      int failed = 0;
      if (threadIdx.x < W && threadIdx.x != r) {
        st_release_sys(tc.pad[threadIdx.x] + flag_index(0, 0, r), e + 1);
        failed = poll_flag(tc.pad[r] + flag_index(0, 0, threadIdx.x), e + 1, tc.timeout_ns, tc.state) ? 0 : 1;
      }
      if (__syncthreads_or(failed)) return;  // a peer is gone: no averaged gradient, no update
      if (threadIdx.x < 208) {
        const float* base = tc.inbox[r] + static_cast<size_t>(e & 1u) * W * 832;
        float4 v[kMaxWorld];
#pragma unroll
        for (int q = 0; q < kMaxWorld; ++q)
          if (q < W) v[q] = ld_f4(base + static_cast<size_t>(q) * 832, threadIdx.x);
        sum = v[0];
#pragma unroll
        for (int q = 1; q < kMaxWorld; ++q)  // fixed rank order: every rank computes the same bits
          if (q < W) { sum.x += v[q].x; sum.y += v[q].y; sum.z += v[q].z; sum.w += v[q].w; }
      }
    } else if (threadIdx.x < 208) {
      sum = reinterpret_cast<const float4*>(row)[threadIdx.x];
    }
    if (threadIdx.x < 208) {
      const int j = threadIdx.x;
      float4 avg = make_float4(sum.x * tc.scale, sum.y * tc.scale, sum.z * tc.scale, sum.w * tc.scale);
      const bool is_w = j < 200;  // float4 200..207 are the 32 bias gradients
      float4* avg_out = is_w ? reinterpret_cast<float4*>(tc.avg_w) + j : reinterpret_cast<float4*>(tc.avg_b) + (j - 200);
      *avg_out = avg;
      if (tc.param_w != nullptr) {
        const SgdHyper h = *tc.hyper;
        const bool has_mom = tc.mom_w != nullptr && h.momentum != 0.0f;
        float4* pp = is_w ? reinterpret_cast<float4*>(tc.param_w) + j : reinterpret_cast<float4*>(tc.param_b) + (j - 200);
        float4* mp = has_mom ? (is_w ? reinterpret_cast<float4*>(tc.mom_w) + j : reinterpret_cast<float4*>(tc.mom_b) + (j - 200)) : nullptr;
        float4 w = *pp, m = has_mom ? *mp : make_float4(0.f, 0.f, 0.f, 0.f);
        w.x = sgd_update(w.x, avg.x, &m.x, h, has_mom);
        w.y = sgd_update(w.y, avg.y, &m.y, h, has_mom);
        w.z = sgd_update(w.z, avg.z, &m.z, h, has_mom);
        w.w = sgd_update(w.w, avg.w, &m.w, h, has_mom);
        *pp = w;
        if (has_mom) *mp = m;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      if (W > 1) tc.state->epoch[0] = e + 1;
      if (tc.param_w != nullptr) tc.hyper->first_step = 0;
    }
  }
}

int c1_imgs_per_cta(int N) { int v = (N + 295) / 296; return v < 1 ? 1 : v; }
int c1_blocks(int N) { int per = c1_imgs_per_cta(N); return (N + per - 1) / per; }

int pool_pix_per_cta(long long total_pix, int C) {
  const int G = kThreads / (C / 4);  // pixel lanes of pool2_bwd_dbias_kernel
  long long p = (total_pix + 511) / 512;  // at most ~512 CTAs (= rows of dbpart)
  if (p < 2 * G) p = 2 * G;
  return static_cast<int>((p + G - 1) / G * G);
}

}  // namespace

// Number of CTAs (= rows of `partial`) pdl_conv1_wgrad_fused uses for a batch of N images.
PDL_API int pdl_bwd_blocks(int N) { return c1_blocks(N); }

// Number of CTAs (= rows of `dbpart`) pdl_pool2_bwd_dbias uses.
PDL_API int pdl_pool2_bwd_blocks(int N, int H, int W, int C) {
  const long long total = static_cast<long long>(N) * H * W;
  const int ppc = pool_pix_per_cta(total, C);
  return static_cast<int>((total + ppc - 1) / ppc);
}

PDL_API int pdl_pool2_bwd_dbias(const float* da, const uint8_t* idx, const float* pooled, float* dz, float* dbpart, int N, int H, int W, int C,
                                int stride, cudaStream_t stream) {
  const long long total = static_cast<long long>(N) * H * W;
  const int ppc = pool_pix_per_cta(total, C);
  const int blocks = static_cast<int>((total + ppc - 1) / ppc);
  if (total * C >= (1ll << 31) || (stride != 1 && stride != 2)) { pdl_set_error("pool2_bwd_dbias: tensor too large or bad stride"); return -1; }
  const unsigned T = static_cast<unsigned>(total);
#define PDL_POOL_CASE(CC, SS) \
  if (C == CC && stride == SS) PDL_CUDA_CHECK(launch_dependent(pool2_bwd_dbias_kernel<CC, SS>, dim3(blocks), dim3(kThreads), 0, stream, da, idx, pooled, dz, dbpart, T, H, W, ppc))
  PDL_POOL_CASE(32, 1); else PDL_POOL_CASE(64, 1); else PDL_POOL_CASE(128, 1); else PDL_POOL_CASE(32, 2); else PDL_POOL_CASE(64, 2);
  else PDL_POOL_CASE(128, 2); else { pdl_set_error("pool2_bwd_dbias: unsupported channel count %d", C); return -1; }
#undef PDL_POOL_CASE
  PDL_LAUNCH_CHECK();
  return 0;
}

// partial must hold pdl_bwd_blocks(N) * 832 floats.
PDL_API int pdl_conv1_wgrad_fused(const float* x, const float* da1, const uint8_t* idx1, const float* a1, float* gw1, float* gb1, float* partial,
                                  int N, cudaStream_t stream) {
  const int smem = (kC1StageFloats + kC1Warps * 832) * static_cast<int>(sizeof(float));
  static bool configured = false;
  if (!configured) {
    PDL_CUDA_CHECK(cudaFuncSetAttribute(conv1_wgrad_fused_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  PDL_CUDA_CHECK(launch_dependent(conv1_wgrad_fused_kernel<false>, dim3(c1_blocks(N)), dim3(kC1Threads), smem, stream, x, da1, idx1, a1, gw1, gb1,
                                  partial, N, c1_imgs_per_cta(N), TailComm{}));
  PDL_LAUNCH_CHECK();
  return 0;
}

// Bytes of symmetric memory the tail exchange needs behind its signal pad: [2][world][832] floats.
PDL_API long long pdl_conv1_tail_inbox_bytes(int world) { return 2ll * world * 832 * static_cast<long long>(sizeof(float)); }

// conv1 weight/bias gradient + cross-GPU average (push one-shot over peer memory) + SGD, ONE kernel.
//   gw1/gb1      raw local gradient (800 / 32 floats)          avg_w/avg_b  averaged gradient out
//   param_*/mom_*/hyper  SGD operands (param_w NULL: average only; mom_w NULL: no momentum)
//   inbox/pad    `world` peer-mapped pointers each (inbox: pdl_conv1_tail_inbox_bytes(world); pad: a signal pad of its own --
//                channel 0 / block 0 of it is used); state: a communicator state used by nothing else.
// world == 1 skips the exchange (inbox/pad may be NULL).  Calls must be stream-ordered per communicator.
PDL_API int pdl_conv1_wgrad_fused_comm(const float* x, const float* da1, const uint8_t* idx1, const float* a1, float* gw1, float* gb1, float* partial,
                                       int N, float* avg_w, float* avg_b, float* param_w, float* param_b, float* mom_w, float* mom_b, void* hyper,
                                       int rank, int world, float* const* inbox, uint32_t* const* pad, void* state, double scale, double timeout_s,
                                       cudaStream_t stream) {
  if (world < 1 || world > pdl_comm::kMaxWorld) { pdl_set_error("conv1 wgrad+allreduce: world %d out of range", world); return -1; }
  if (world > 1 && (!inbox || !pad || !state)) { pdl_set_error("conv1 wgrad+allreduce: peer pointers missing"); return -1; }
  if (param_w && !hyper) { pdl_set_error("conv1 wgrad+allreduce: SGD needs hyper-parameters"); return -1; }
  const int smem = (kC1StageFloats + kC1Warps * 832) * static_cast<int>(sizeof(float));
  static bool configured = false;
  if (!configured) {
    PDL_CUDA_CHECK(cudaFuncSetAttribute(conv1_wgrad_fused_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  TailComm tc{};
  for (int q = 0; q < world && world > 1; ++q) { tc.inbox[q] = inbox[q]; tc.pad[q] = pad[q]; }
  tc.state = static_cast<pdl_comm::CommState*>(state);
  tc.avg_w = avg_w; tc.avg_b = avg_b; tc.param_w = param_w; tc.param_b = param_b; tc.mom_w = mom_w; tc.mom_b = mom_b;
  tc.hyper = static_cast<SgdHyper*>(hyper); tc.rank = rank; tc.world = world; tc.scale = static_cast<float>(scale);
  tc.timeout_ns = static_cast<unsigned long long>(timeout_s * 1e9);
  PDL_CUDA_CHECK(launch_dependent(conv1_wgrad_fused_kernel<true>, dim3(c1_blocks(N)), dim3(kC1Threads), smem, stream, x, da1, idx1, a1, gw1, gb1,
                                  partial, N, c1_imgs_per_cta(N), tc));
  PDL_LAUNCH_CHECK();
  return 0;
}
