// Generic fp32 implicit-GEMM tile kernel on CUDA cores (FFMA), used as the any-shape fallback
// and as the in-library oracle for the tcgen05 kernels.  C[M,N] = sum_k A(m,k) * B(k,n) where
// A, B and the C store are functors, so convolution forward / dgrad / wgrad are three small
// index mappings over the same mainloop.  Strict fp32 (no TF32), deterministic.
#pragma once
#include "common.cuh"

namespace simt {

constexpr int BM = 64, BN = 64, BK = 16, THREADS = 256;

template <typename ALoad, typename BLoad, typename CStore>
__global__ void __launch_bounds__(THREADS) igemm_kernel(int M, int N, int K, int k_per_split, ALoad a_load, BLoad b_load, CStore c_store) {
  __shared__ float As[BK][BM + 4];
  __shared__ float Bs[BK][BN + 4];
  const int tid = threadIdx.x;
  const int m0 = blockIdx.x * BM, n0 = blockIdx.y * BN;
  const int k_begin = blockIdx.z * k_per_split;
  const int k_end = min(K, k_begin + k_per_split);
  const int tx = tid % 16, ty = tid / 16;  // 16x16 threads, 4x4 outputs each
  float acc[4][4] = {};

  for (int k0 = k_begin; k0 < k_end; k0 += BK) {
    // A tile: BM x BK, element (m, k); consecutive threads walk m so NHWC/row gathers coalesce per functor
#pragma unroll
    for (int i = 0; i < (BM * BK) / THREADS; ++i) {
      int e = tid + i * THREADS;
      int kk = e / BM, mm = e % BM;
      int m = m0 + mm, k = k0 + kk;
      As[kk][mm] = (m < M && k < k_end) ? a_load(m, k) : 0.0f;
    }
#pragma unroll
    for (int i = 0; i < (BN * BK) / THREADS; ++i) {
      int e = tid + i * THREADS;
      int kk = e / BN, nn = e % BN;
      int n = n0 + nn, k = k0 + kk;
      Bs[kk][nn] = (n < N && k < k_end) ? b_load(k, n) : 0.0f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < BK; ++kk) {
      float a[4], b[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) a[i] = As[kk][ty * 4 + i];
#pragma unroll
      for (int j = 0; j < 4; ++j) b[j] = Bs[kk][tx * 4 + j];
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], b[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      int m = m0 + ty * 4 + i, n = n0 + tx * 4 + j;
      if (m < M && n < N) c_store(m, n, acc[i][j], static_cast<int>(blockIdx.z));
    }
}

template <typename ALoad, typename BLoad, typename CStore>
int launch(int M, int N, int K, int splits, ALoad a, BLoad b, CStore c, cudaStream_t stream) {
  if (M <= 0 || N <= 0) return 0;
  if (splits < 1) splits = 1;
  int k_per_split = ((K + splits - 1) / splits + BK - 1) / BK * BK;
  splits = (K + k_per_split - 1) / k_per_split;
  if (splits < 1) splits = 1;
  dim3 grid((M + BM - 1) / BM, (N + BN - 1) / BN, splits);
  igemm_kernel<<<grid, THREADS, 0, stream>>>(M, N, K, k_per_split, a, b, c);
  PDL_LAUNCH_CHECK();
  return 0;
}

inline int num_splits(int K, int splits) {
  if (splits < 1) splits = 1;
  int k_per_split = ((K + splits - 1) / splits + BK - 1) / BK * BK;
  int s = (K + k_per_split - 1) / k_per_split;
  return s < 1 ? 1 : s;
}

}  // namespace simt
