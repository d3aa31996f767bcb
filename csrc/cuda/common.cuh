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
