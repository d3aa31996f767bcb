// Shared helpers for the sm_100a kernel library (plain C ABI, loaded with ctypes).
#pragma once
#include <cuda_runtime.h>
#include <cstdint>
#include <cstdio>

#define PDL_API extern "C" __attribute__((visibility("default")))

// Every entry point returns 0 on success or a negative/CUDA error code; the message is kept
// in a thread-local buffer readable through pdl_last_error().
void pdl_set_error(const char* fmt, ...);

#define PDL_CUDA_CHECK(expr)                                                                  \
  do {                                                                                        \
    cudaError_t _e = (expr);                                                                  \
    if (_e != cudaSuccess) {                                                                  \
      pdl_set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__, cudaGetErrorString(_e)); \
      return static_cast<int>(_e);                                                            \
    }                                                                                         \
  } while (0)

// Every launcher ends with this: counts the call (pdl_launch_count(), used to report how many of this
// library's launchers ran in a timed region) and surfaces launch errors.
extern long long g_pdl_launch_calls;
#define PDL_LAUNCH_CHECK() do { ++g_pdl_launch_calls; PDL_CUDA_CHECK(cudaGetLastError()); } while (0)

static inline int pdl_ceil_div(long long a, long long b) { return static_cast<int>((a + b - 1) / b); }

// ---- programmatic dependent launch ----
// The training step is a chain of short kernels, each a fraction of one wave: what separates them is launch +
// drain latency.  A kernel launched through launch_dependent() may start while its predecessor in the stream is
// still draining and run its prologue (shared-memory carve-up, barrier init, TMEM allocation, descriptor prefetch);
// it must call grid_dependency_wait() before it touches anything the predecessor wrote (the wait returns once the
// predecessor grid has completed and its writes are visible; it is a no-op for ordinary launches).  Kernels call
// launch_dependents() first thing so that the successor can be scheduled as soon as all of their CTAs are resident.
// PDL_NO_PDL=1 turns the launch attribute off (plain stream order).
__device__ __forceinline__ void grid_dependency_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_dependent_launch_enabled();

template <typename... KArgs, typename... Args>
cudaError_t launch_dependent(void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg{};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = pdl_dependent_launch_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}

// Round-to-nearest TF32 (10-bit mantissa).  The tensor cores TRUNCATE fp32 operands to TF32; producers of
// MMA operands round on store instead, which halves the error and removes the toward-zero bias.
__device__ __forceinline__ float round_tf32(float x) {
  uint32_t u;
  asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(u) : "f"(x));
  return __uint_as_float(u);
}

struct SgdHyper {  // lives in device memory so a captured CUDA graph sees LR-schedule changes
  float lr;
  float momentum;
  float dampening;
  float weight_decay;
  int nesterov;
  int first_step;  // 1 until the first update has run (momentum buffer := gradient)
  int pad0, pad1;
};

// One SGD update, identical to the rule the reference's optimizer applies
// (launch_dist.py:59 plain SGD; example_launch.py:64-70 momentum+wd+nesterov).
__device__ __forceinline__ float sgd_update(float p, float g, float* m_inout, const SgdHyper& h, bool has_momentum) {
  g = fmaf(h.weight_decay, p, g);
  if (has_momentum) {
    float buf = h.first_step ? g : fmaf(h.momentum, *m_inout, (1.0f - h.dampening) * g);
    *m_inout = buf;
    g = h.nesterov ? fmaf(h.momentum, buf, g) : buf;
  }
  return fmaf(-h.lr, g, p);
}
