// Cross-GPU signalling primitives shared by the collective kernels (allreduce.cu) and by compute kernels that finish
// with a collective of their own (convnet_bwd.cu: conv1 weight gradient + tail allreduce + SGD in one kernel).
// Flags are monotone epochs in symmetric memory: written with st.release.sys into the PEER's pad, polled with
// ld.acquire.sys in the LOCAL pad.  Every wait is bounded; a timeout flags the communicator (device word + optional
// pinned host word) and callers must then leave gradients and parameters untouched.
#pragma once
#include "common.cuh"

namespace pdl_comm {

constexpr int kMaxWorld = 8;
constexpr int kMaxBlocks = 128;
constexpr int kMaxChannels = 8;

// Signal pad: flag[channel][block][source rank].  Lives at the start of a symmetric segment.
__host__ __device__ inline size_t flag_index(int ch, int block, int src) { return (static_cast<size_t>(ch) * kMaxBlocks + block) * kMaxWorld + src; }

struct CommState {  // local (non-symmetric) device memory, one per communicator
  uint32_t epoch[kMaxChannels];
  uint32_t finished[kMaxChannels];
  uint32_t error;        // sticky: set to 1 when a peer wait timed out; every later launch on this communicator is a no-op
  uint32_t reserved;
  uint32_t* host_error;  // optional pinned host word (device-visible): lets the host notice without a D2H copy or a sync
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) { asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory"); }
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long globaltimer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
__device__ __forceinline__ float4 ld_f4(const float* base, long long i) {
  float4 v;
  asm volatile("ld.global.relaxed.sys.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(base + 4 * i) : "memory");
  return v;
}
__device__ __forceinline__ void st_f4(float* base, long long i, float4 v) {
  asm volatile("st.global.relaxed.sys.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(base + 4 * i), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void flag_failure(CommState* state) {
  state->error = 1;
  uint32_t* host = state->host_error;
  if (host) *host = 1;
  __threadfence_system();
}
// One thread polls one flag until it reaches `v` (epochs wrap: signed distance).  false = timed out (communicator flagged).
__device__ __forceinline__ bool poll_flag(const uint32_t* f, uint32_t v, unsigned long long timeout_ns, CommState* state) {
  unsigned long long t0 = 0;
  uint32_t polls = 0;
  while (static_cast<int32_t>(ld_acquire_sys(f) - v) < 0) {
    if ((++polls & 1023u) == 0) {  // consult the clock rarely: the poll itself is the hot path
      const unsigned long long now = globaltimer_ns();
      if (t0 == 0) t0 = now;
      else if (now - t0 > timeout_ns) { flag_failure(state); return false; }
    }
  }
  return true;
}

}  // namespace pdl_comm
