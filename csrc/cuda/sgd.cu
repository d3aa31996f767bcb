// Stand-alone SGD update: one launch for a whole parameter list (multi-tensor) or a flat arena.
// Used when the update is NOT fused into the allreduce kernel: world_size 1 without DDP,
// no_sync iterations, parameters outside any bucket.  Update rule = reference optimizer
// (launch_dist.py:59, example_launch.py:64-70), see sgd_update() in common.cuh.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kChunk = 4096;  // elements per CTA

struct TensorRef {  // one row of the device-resident table
  float* p;
  const float* g;
  float* m;  // nullptr when momentum == 0
  long long n;
};

__global__ void __launch_bounds__(kThreads) multi_tensor_sgd_kernel(const TensorRef* __restrict__ table, const int* __restrict__ chunk_prefix,
                                                                    int num_tensors, SgdHyper h) {
  // binary search: which tensor does this chunk belong to
  int lo = 0, hi = num_tensors;
  const int c = blockIdx.x;
  while (hi - lo > 1) {
    int mid = (lo + hi) >> 1;
    if (chunk_prefix[mid] <= c) lo = mid; else hi = mid;
  }
  const TensorRef t = table[lo];
  const long long base = static_cast<long long>(c - chunk_prefix[lo]) * kChunk;
  const long long end = min(t.n, base + kChunk);
  const bool has_mom = t.m != nullptr && h.momentum != 0.0f;
  const bool vec = ((reinterpret_cast<uintptr_t>(t.p) | reinterpret_cast<uintptr_t>(t.g) | reinterpret_cast<uintptr_t>(t.m)) & 15) == 0;
  if (vec) {
    long long i = base + threadIdx.x * 4;
    for (; i + 3 < end; i += kThreads * 4) {
      float4 w = *reinterpret_cast<float4*>(t.p + i);
      float4 g = *reinterpret_cast<const float4*>(t.g + i);
      float4 m = has_mom ? *reinterpret_cast<float4*>(t.m + i) : make_float4(0, 0, 0, 0);
      w.x = sgd_update(w.x, g.x, &m.x, h, has_mom);
      w.y = sgd_update(w.y, g.y, &m.y, h, has_mom);
      w.z = sgd_update(w.z, g.z, &m.z, h, has_mom);
      w.w = sgd_update(w.w, g.w, &m.w, h, has_mom);
      *reinterpret_cast<float4*>(t.p + i) = w;
      if (has_mom) *reinterpret_cast<float4*>(t.m + i) = m;
    }
    // tail (n not a multiple of 4): the thread whose vector would cross `end`
    if (i < end) {
      for (long long j = i; j < end; ++j) {
        float m = has_mom ? t.m[j] : 0.0f;
        t.p[j] = sgd_update(t.p[j], t.g[j], &m, h, has_mom);
        if (has_mom) t.m[j] = m;
      }
    }
  } else {
    for (long long i = base + threadIdx.x; i < end; i += kThreads) {
      float m = has_mom ? t.m[i] : 0.0f;
      t.p[i] = sgd_update(t.p[i], t.g[i], &m, h, has_mom);
      if (has_mom) t.m[i] = m;
    }
  }
}

}  // namespace

PDL_API int pdl_sgd_chunk_elems() { return kChunk; }

// table: device array of num_tensors rows {p, g, m, n}; chunk_prefix: device int[num_tensors+1].
PDL_API int pdl_multi_tensor_sgd(const void* table, const int* chunk_prefix, int num_tensors, int total_chunks, float lr, float momentum,
                                 float dampening, float weight_decay, int nesterov, int first_step, cudaStream_t stream) {
  if (num_tensors == 0 || total_chunks == 0) return 0;
  SgdHyper h{lr, momentum, dampening, weight_decay, nesterov, first_step, 0, 0};
  multi_tensor_sgd_kernel<<<total_chunks, kThreads, 0, stream>>>(static_cast<const TensorRef*>(table), chunk_prefix, num_tensors, h);
  PDL_LAUNCH_CHECK();
  return 0;
}
