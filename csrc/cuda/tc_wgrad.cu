// Convolution weight gradient on tcgen05 tensor cores (TF32 in, FP32 accumulate in TMEM).
//
//   dW[co][ci][r][s] = sum over pixels (n,p,q) of dy[n,p,q,co] * x[n, p+r-pad, q+s-pad, ci]
//
// GEMM view: M = Cout, N = Cin (one accumulator per filter tap), K = pixels.  Both operands are
// "MN-major" (channels contiguous, one 128-byte row per pixel), which is exactly how TMA lands an
// NHWC box in shared memory -- no transposes, and the filter tap only shifts the x box in the outer
// (W, H) coordinates.  32-bit MN-major operands need the 32-byte-atom flavour of the 128-byte
// swizzle on both sides (TMA: SWIZZLE_128B_ATOM_32B, UMMA descriptor: SWIZZLE_128B_BASE32B, 4-row
// K atoms); the plain SWIZZLE_128B descriptor silently yields zeros for TF32 (measured).  A pixel
// box is (32 ch, BW, BH, 1) with BW rounded up to a multiple of 8; TMA zero-fills the out-of-range
// columns/rows of dy, so padded rows add exact zeros.  The dy tile is reused for the taps of one
// filter row, K is split over CTAs, each CTA writes partial sums to a scratch buffer, and a second
// kernel reduces the splits in fixed order (deterministic) into the gradient bucket and finishes the
// bias gradient.  Replaces cuDNN wgrad + ATen bias reductions.
#include "comm.cuh"
#include "tc_common.cuh"

#include <cstdlib>

namespace {

constexpr int kThreads = 256;
constexpr int kMaxTapsPerCta = 7;  // 3 taps of one filter row (generic), or all 7 rows of a small-Cin "stem" filter

struct WgGeom {
  int BW, BH;            // pixel box (BW % 8 == 0), rows = BW*BH in {64, 128}
  int strips, wstrips;   // row strips / column strips per image
  int units;             // N * strips * wstrips  (K work items)
  int units_per_cta;     // K split
  int S, RT;             // taps per filter row x filter rows handled by one CTA (S * RT accumulators)
  int pad, stride;       // stride applies to the row coordinate of x
  int stride_w;          // multiplier of the column coordinate of x (1 in the stem's virtual layout)
  int layout_type, lbo, sbo, tma_swizzle;  // descriptor knobs (defaults set on the host)
  int Cin, Cout, cin_tiles;  // full channel counts; blockIdx.z = cout_tile * cin_tiles + cin_tile
};

__device__ __forceinline__ uint64_t smem_desc_lt(uint32_t smem_addr, uint32_t lbo, uint32_t sbo, uint32_t layout_type) {
  uint64_t d = 0;
  d |= static_cast<uint64_t>((smem_addr >> 4) & 0x3FFF);
  d |= static_cast<uint64_t>((lbo >> 4) & 0x3FFF) << 16;
  d |= static_cast<uint64_t>((sbo >> 4) & 0x3FFF) << 32;
  d |= static_cast<uint64_t>(1) << 46;
  d |= static_cast<uint64_t>(layout_type & 7) << 61;
  return d;
}

template <int M_GEMM, int N_GEMM>  // M_GEMM = Cout (64|128), N_GEMM = Cin (32|64)
__global__ void __launch_bounds__(kThreads, 1)
tc_wgrad_kernel(const __grid_constant__ CUtensorMap tmDy, const __grid_constant__ CUtensorMap tmX, const WgGeom g, float* __restrict__ partial,
                int* __restrict__ error) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  const int rows = g.BW * g.BH;
  const int chunk_bytes = rows * 128;
  constexpr int A_CHUNKS = M_GEMM / 32, B_CHUNKS = N_GEMM / 32;
  const int a_bytes = A_CHUNKS * chunk_bytes;
  const int b_tap_bytes = B_CHUNKS * chunk_bytes;
  const int T = g.S * g.RT;  // accumulators (filter taps) of this CTA
  const int stage_bytes = a_bytes + T * b_tap_bytes;
  const int stages = (stage_bytes * 4 <= 200 * 1024) ? 4 : (stage_bytes * 3 <= 200 * 1024) ? 3 : 2;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + stages * stage_bytes);
  uint64_t* empty = full + 4;
  uint64_t* done = empty + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(done + 1);
  volatile int* failed = reinterpret_cast<volatile int*>(tmem_slot + 1);

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
  const uint32_t need_cols = static_cast<uint32_t>(T * N_GEMM);
  const uint32_t kCols = need_cols <= 32 ? 32 : need_cols <= 64 ? 64 : need_cols <= 128 ? 128 : need_cols <= 256 ? 256 : 512;
  const int tap_row = blockIdx.x;  // filter row r
  const int split = blockIdx.y;
  const int co0 = (blockIdx.z / g.cin_tiles) * M_GEMM, ci0 = (blockIdx.z % g.cin_tiles) * N_GEMM;  // channel tile origin
  const int u_begin = split * g.units_per_cta;
  const int u_end = min(g.units, u_begin + g.units_per_cta);

  if (threadIdx.x == 0) {
    tc::prefetch_tmap(&tmDy);
    tc::prefetch_tmap(&tmX);
    for (int i = 0; i < 4; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], 1); }
    tc::mbar_init(done, 1);
    *failed = 0;
    tc::fence_barrier_init();
  }
  if (warp == 2) tc::tmem_alloc(tmem_slot, kCols);
  launch_dependents();     // successor may set up while this grid drains
  grid_dependency_wait();  // predecessor's outputs are complete and visible from here on
  tc::tc_fence_before();
  __syncthreads();
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    if (tc::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      for (int u = u_begin; u < u_end; ++u) {
        if (!tc::mbar_wait(&empty[stage], phase ^ 1)) { *failed = 1; break; }
        const int per_img = g.strips * g.wstrips;
        const int n = u / per_img, rem = u % per_img;
        const int h0 = (rem / g.wstrips) * g.BH, w0 = (rem % g.wstrips) * g.BW;
        uint8_t* base = smem + stage * stage_bytes;
        tc::mbar_expect_tx(&full[stage], static_cast<uint32_t>(stage_bytes));
        for (int c = 0; c < A_CHUNKS; ++c) tc::tma_load_4d(base + c * chunk_bytes, &tmDy, &full[stage], co0 + c * 32, w0, h0, n);
        for (int t = 0; t < T; ++t) {
          const int r = tap_row * g.RT + t / g.S, sx = t % g.S;
          for (int c = 0; c < B_CHUNKS; ++c)
            tc::tma_load_4d(base + a_bytes + t * b_tap_bytes + c * chunk_bytes, &tmX, &full[stage], ci0 + c * 32, w0 * g.stride_w + sx - g.pad,
                            h0 * g.stride + r - g.pad, n);
        }
        if (++stage == stages) { stage = 0; phase ^= 1; }
      }
    }
  } else if (warp == 1) {
    constexpr uint32_t idesc = tc::idesc_tf32(M_GEMM, N_GEMM, 1, 1);  // both operands MN-major
    const uint32_t lbo = g.lbo ? g.lbo : chunk_bytes, sbo = g.sbo;
    int stage = 0;
    uint32_t phase = 0;
    for (int u = u_begin; u < u_end; ++u) {
      if (!tc::mbar_wait(&full[stage], phase)) { *failed = 1; break; }
      tc::tc_fence_after();
      if (tc::elect_one()) {
        const uint32_t a_addr = tc::smem_u32(smem + stage * stage_bytes);
        const uint32_t b_addr = a_addr + a_bytes;
        for (int k8 = 0; k8 < rows / 8; ++k8) {  // one MMA = 8 pixels = 8 rows of 128 B
          const uint64_t a_desc = smem_desc_lt(a_addr + k8 * 1024, lbo, sbo, g.layout_type);
          for (int t = 0; t < T; ++t) {
            const uint64_t b_desc = smem_desc_lt(b_addr + t * b_tap_bytes + k8 * 1024, lbo, sbo, g.layout_type);
            tc::mma_tf32(tmem_base + t * N_GEMM, a_desc, b_desc, idesc, (u != u_begin) || (k8 != 0));
          }
        }
        tc::mma_commit(&empty[stage]);
        if (u == u_end - 1) tc::mma_commit(done);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1; }
    }
  } else if (warp >= 4) {
    const int quarter = warp & 3;
    const bool has_work = u_end > u_begin;
    bool ok = true;
    if (has_work) ok = tc::mbar_wait(done, 0);
    if (!ok) *failed = 1;
    tc::tc_fence_after();
    // accumulator row -> output channel: M=128: lane index; M=64: 16 rows per 32-lane quarter
    const int co = (M_GEMM == 128) ? quarter * 32 + lane : quarter * 16 + lane;
    const bool row_valid = (M_GEMM == 128) || lane < 16;
    for (int s = 0; s < T; ++s) {
      const int tap = tap_row * T + s;
      float* dst = partial + ((static_cast<long long>(split) * (gridDim.x * T) + tap) * g.Cout + co0 + co) * g.Cin + ci0;
#pragma unroll 1
      for (int c = 0; c < N_GEMM; c += 32) {
        float v[32];
        if (has_work && ok) {
          tc::tmem_ld32(tmem_base + s * N_GEMM + c + (static_cast<uint32_t>(quarter * 32) << 16), v);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0.f;
        }
        if (row_valid) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) *reinterpret_cast<float4*>(dst + c + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
      }
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (warp == 2) tc::tmem_dealloc(tmem_base, kCols);
  if (threadIdx.x == 0 && *failed && error) *error = 1;
}

// dw[co][ci][tap] = sum_split partial[split][tap][co][ci];  db[co] = sum_blk db_partial[blk][co]
// One thread per OUTPUT element, gathering its `splits` partials (stride Cout*Cin floats between neighbouring threads).  Two
// "coalesced" rewrites -- a CTA per output-channel block, and a CTA per (channels x 32 input channels x taps) tile, each
// reading 128-byte rows and transposing through shared memory -- were measured SLOWER in round 2 (ResNet-50 step 14.25 and
// 13.38 ms against 13.08 ms; ConvNet 111-115 against 101 us): the kernel is bound by load latency, not by sectors, and the
// gather keeps one independent chain per output with no barrier in it.
__global__ void wgrad_finish_kernel(const float* __restrict__ partial, int splits, int taps, int Cout, int Cin, float* __restrict__ dw,
                                    const float* __restrict__ db_partial, int db_blocks, float* __restrict__ db) {
  launch_dependents();     // successor may set up while this grid drains
  grid_dependency_wait();  // predecessor's outputs are complete and visible from here on
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  const int per_split = taps * Cout * Cin;
  if (i < per_split) {
    // i indexes the OUTPUT layout so writes are coalesced
    const int tap = i % taps, ci = (i / taps) % Cin, co = i / (taps * Cin);
    const float* src = partial + (static_cast<long long>(tap) * Cout + co) * Cin + ci;
    float acc = 0.f;
#pragma unroll 8
    for (int s = 0; s < splits; ++s) acc += __ldg(src + static_cast<long long>(s) * per_split);  // independent loads, fixed order
    dw[i] = acc;
  }
  // bias gradient: one warp per channel, lanes stride over the partial rows, fixed shuffle tree (deterministic)
  if (db) {
    const int warp_global = (blockIdx.x * blockDim.x + threadIdx.x) / 32, lane = threadIdx.x % 32;
    const int first = (per_split + 31) / 32;  // warps past the weight outputs
    const int co = warp_global - first;
    if (co >= 0 && co < Cout) {
      float acc = 0.f;
      for (int b = lane; b < db_blocks; b += 32) acc += db_partial[b * Cout + co];
#pragma unroll
      for (int o = 16; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
      if (lane == 0) db[co] = acc;
    }
  }
}

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn encode_tiled_wg() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}
int act_map(CUtensorMap* m, const float* base, int N, int H, int W, int C, int bw, int bh, int swizzle, int stride) {
  EncodeTiledFn enc = encode_tiled_wg();
  if (!enc) { pdl_set_error("cuTensorMapEncodeTiled is not available"); return -1; }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
  cuuint32_t box[4] = {32, (cuuint32_t)(bw * stride), (cuuint32_t)(bh * stride), 1};  // traversal stride keeps bw x bh pixels
  cuuint32_t es[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   static_cast<CUtensorMapSwizzle>(swizzle), CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { pdl_set_error("cuTensorMapEncodeTiled(wgrad operand) failed: %d", (int)r); return -2; }
  return 0;
}

int env_int(const char* name, int dflt) { const char* v = getenv(name); return v ? atoi(v) : dflt; }

// Upper bound on the CTAs of the next weight-gradient launches (0 = none); see pdl_tc_wgrad_max_ctas.
int g_wg_cta_cap = 0;

// Channel tiles: M (= Cout) tile 128 (64 when Cout == 64).  N (= Cin) tile as wide as the TMEM budget allows --
// every operand byte comes through L2 (~42 B/clk/SM), so FLOPs per loaded byte is what bounds this kernel:
// 1x1 filters take 256 / 128 columns, filter rows of up to 3 taps 128 (3 accumulators = 384 TMEM columns).
int wg_m_tile(int Cout) { return Cout == 64 ? 64 : (Cout % 128 == 0 ? 128 : 0); }
int wg_n_tile(int Cin, int KW = 3) {
  if (Cin == 32) return 32;
  if (KW == 1 && Cin % 256 == 0) return 256;
  if (Cin % 128 == 0) return 128;
  return Cin % 64 == 0 ? 64 : 0;
}

bool wg_geometry(int N, int P, int Q, int Cin, int Cout, int KH, int KW, int pad, WgGeom* g) {
  const int mt = wg_m_tile(Cout), nt = wg_n_tile(Cin, KW);
  if (!mt || !nt || KW > 3) return false;
  int bw = (Q + 7) / 8 * 8;  // K stage = bw x bh pixels, a multiple of 8 rows (one MMA = 8 pixels)
  if (bw > 64) return false;
  // wide channel tiles take 32-pixel stages so that 3-4 stages still fit in shared memory
  const int stage_rows = (mt / 32 + KW * (nt / 32)) > 10 ? 32 : 64;
  int bh = stage_rows / bw;
  if (bh < 1) bh = 1;
  g->BW = bw; g->BH = bh; g->strips = (P + bh - 1) / bh; g->wstrips = 1; g->units = N * g->strips;
  g->RT = 1; g->stride_w = 1;
  const int rows = bw * bh;
  const long long stage_bytes = static_cast<long long>(mt / 32 + KW * (nt / 32)) * rows * 128;
  if (stage_bytes * 2 > 200 * 1024) return false;
  g->Cin = Cin; g->Cout = Cout; g->cin_tiles = Cin / nt;
  // K split: enough CTAs to occupy the chip (grid = KH x splits x channel tiles), few enough that the partial sums stay cheap
  const int tiles = (Cout / mt) * (Cin / nt);
  int splits = (2 * 148 + KH * tiles - 1) / (KH * tiles);
  if (KH * tiles < 74) splits = 148 / (KH * tiles);  // few big tiles: one wave, each split costs a full dW of partial traffic
  if (splits > 148) splits = 148;
  {
    static const int env_cap = env_int("PDL_WG_MAX_CTAS", 0);  // experiment knob: bound the grid (KH * splits * tiles)
    const int cta_cap = env_cap > 0 ? env_cap : g_wg_cta_cap;
    if (cta_cap > 0 && KH * tiles * splits > cta_cap) splits = cta_cap / (KH * tiles) > 0 ? cta_cap / (KH * tiles) : 1;
  }
  if (splits > g->units) splits = g->units;
  if (splits < 1) splits = 1;
  g->units_per_cta = (g->units + splits - 1) / splits;
  g->S = KW; g->pad = pad; g->stride = 1;
  g->layout_type = env_int("PDL_WG_LT", 1);   // SWIZZLE_128B_BASE32B
  g->lbo = env_int("PDL_WG_LBO", 0);          // 0 = distance between 32-channel chunks
  g->sbo = env_int("PDL_WG_SBO", 512);        // 4-row K atoms, dense
  g->tma_swizzle = env_int("PDL_WG_TMASW", static_cast<int>(CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B));
  return true;
}
int wg_splits(const WgGeom& g) { return (g.units + g.units_per_cta - 1) / g.units_per_cta; }

template <int M, int Nn>
int launch_wg(const CUtensorMap& a, const CUtensorMap& b, const WgGeom& g, int KH, float* partial, int* error, cudaStream_t stream) {
  auto kernel = tc_wgrad_kernel<M, Nn>;
  static bool configured = false;
  const int smem = 210 * 1024;
  if (!configured) {
    PDL_CUDA_CHECK(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid(KH, wg_splits(g), (g.Cout / M) * g.cin_tiles);
  PDL_CUDA_CHECK(launch_dependent(kernel, grid, dim3(kThreads), smem, stream, a, b, g, partial, error));
  PDL_LAUNCH_CHECK();
  return 0;
}

}  // namespace

// Bound the grid of the weight-gradient launches that follow (0 lifts the bound); returns the previous value.  A caller
// that runs a small weight gradient NEXT TO other kernels uses this: every CTA of these kernels takes a whole SM's shared
// memory, so a one-wave grid makes concurrent kernels queue for SMs, while a third of a wave leaves room for them (and
// leaves the second pass a third of the partial sums).  Measured on the batch-100 ConvNet step: 124.9 us with the one-wave
// grid, 116.9 us with 48 CTAs, 123-124 us at 36 or 60.  Host-side setting, read when a launch is issued (or captured).
PDL_API int pdl_tc_wgrad_max_ctas(int cap) {
  const int old = g_wg_cta_cap;
  g_wg_cta_cap = cap < 0 ? 0 : cap;
  return old;
}

PDL_API int pdl_tc_wgrad_supported(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad) {
  if (stride != 1 && stride != 2) return 0;
  WgGeom g{};
  return wg_geometry(N, (H + 2 * pad - KH) / stride + 1, (W + 2 * pad - KW) / stride + 1, Cin, Cout, KH, KW, pad, &g) ? 1 : 0;
}

PDL_API long long pdl_tc_wgrad_scratch_bytes_strided(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad) {
  WgGeom g{};
  if (!wg_geometry(N, (H + 2 * pad - KH) / stride + 1, (W + 2 * pad - KW) / stride + 1, Cin, Cout, KH, KW, pad, &g)) return 0;
  return static_cast<long long>(wg_splits(g)) * KH * KW * Cout * Cin * sizeof(float);
}
PDL_API long long pdl_tc_wgrad_scratch_bytes(int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad) {
  return pdl_tc_wgrad_scratch_bytes_strided(N, H, W, Cin, Cout, KH, KW, 1, pad);
}

// x[N][H][W][Cin], dy[N][P][Q][Cout] (NHWC) -> dw[Cout][Cin][KH][KW], convolution stride 1 or 2
// (+ db[Cout] = fixed-order sum of db_partial[db_blocks][Cout] when db != nullptr).
PDL_API int pdl_tc_conv2d_wgrad_strided(const float* x, const float* dy, float* dw, float* db, const float* db_partial, int db_blocks,
                                        float* scratch, int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad,
                                        int* error_flag, cudaStream_t stream);
PDL_API int pdl_tc_conv2d_wgrad(const float* x, const float* dy, float* dw, float* db, const float* db_partial, int db_blocks, float* scratch,
                                int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad, int* error_flag, cudaStream_t stream) {
  return pdl_tc_conv2d_wgrad_strided(x, dy, dw, db, db_partial, db_blocks, scratch, N, H, W, Cin, Cout, KH, KW, 1, pad, error_flag, stream);
}
PDL_API int pdl_tc_conv2d_wgrad_strided(const float* x, const float* dy, float* dw, float* db, const float* db_partial, int db_blocks,
                                        float* scratch, int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad,
                                        int* error_flag, cudaStream_t stream) {
  WgGeom g{};
  if (stride != 1 && stride != 2) { pdl_set_error("tc wgrad: stride %d is not supported", stride); return -3; }
  const int P = (H + 2 * pad - KH) / stride + 1, Q = (W + 2 * pad - KW) / stride + 1;
  if (!wg_geometry(N, P, Q, Cin, Cout, KH, KW, pad, &g)) { pdl_set_error("tc wgrad: unsupported shape"); return -3; }
  g.stride = g.stride_w = stride;
  CUtensorMap ta, tb;
  if (int rc = act_map(&ta, dy, N, P, Q, Cout, g.BW, g.BH, g.tma_swizzle, 1)) return rc;
  if (int rc = act_map(&tb, x, N, H, W, Cin, g.BW, g.BH, g.tma_swizzle, stride)) return rc;
  int rc = -3;
  const int mt = wg_m_tile(Cout), nt = wg_n_tile(Cin, KW);
  if (mt == 128 && nt == 256) rc = launch_wg<128, 256>(ta, tb, g, KH, scratch, error_flag, stream);
  else if (mt == 128 && nt == 128) rc = launch_wg<128, 128>(ta, tb, g, KH, scratch, error_flag, stream);
  else if (mt == 64 && nt == 256) rc = launch_wg<64, 256>(ta, tb, g, KH, scratch, error_flag, stream);
  else if (mt == 64 && nt == 128) rc = launch_wg<64, 128>(ta, tb, g, KH, scratch, error_flag, stream);
  else if (mt == 128 && nt == 64) rc = launch_wg<128, 64>(ta, tb, g, KH, scratch, error_flag, stream);
  else if (mt == 128 && nt == 32) rc = launch_wg<128, 32>(ta, tb, g, KH, scratch, error_flag, stream);
  else if (mt == 64 && nt == 32) rc = launch_wg<64, 32>(ta, tb, g, KH, scratch, error_flag, stream);
  else if (mt == 64 && nt == 64) rc = launch_wg<64, 64>(ta, tb, g, KH, scratch, error_flag, stream);
  if (rc) return rc;
  if (dw == nullptr) return 0;  // GEMM pass only (pdl_tc_conv2d_wgrad_gemm); the second pass is launched separately
  const int total = (KH * KW * Cout * Cin + 31) / 32 * 32 + (db ? Cout * 32 : 0);
  PDL_CUDA_CHECK(launch_dependent(wgrad_finish_kernel, dim3(pdl_ceil_div(total, 256)), dim3(256), 0, stream, scratch, wg_splits(g), KH * KW, Cout, Cin, dw,
                                  db_partial, db_blocks, db));
  PDL_LAUNCH_CHECK();
  return 0;
}

// ---------------------------------------------------------------------------------------------------------
// Small-Cin ("stem") weight gradient: same virtual layout as the forward (tc::stem_dims).  GEMM per filter
// row r: M = Cout, N = 32 (KW taps x 4 channels), K = pixels; one CTA accumulates all KH rows of its K slice.
// ---------------------------------------------------------------------------------------------------------
namespace {

// dw[co][c][r][s] = sum_split partial[split][r][co][s*4 + c]
__global__ void stem_wgrad_finish_kernel(const float* __restrict__ partial, int splits, int Cout, int Cin, int KH, int KW, float* __restrict__ dw) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= Cout * Cin * KH * KW) return;
  const int s = i % KW, r = (i / KW) % KH, c = (i / (KW * KH)) % Cin, co = i / (KW * KH * Cin);
  const float* src = partial + (static_cast<long long>(r) * Cout + co) * 32 + s * 4 + c;
  const long long per_split = static_cast<long long>(KH) * Cout * 32;
  float acc = 0.f;
#pragma unroll 8
  for (int k = 0; k < splits; ++k) acc += __ldg(src + k * per_split);
  dw[i] = acc;
}

bool stem_wg_geometry(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad, WgGeom* g) {
  const int mt = wg_m_tile(Cout);
  if (!mt || Cin < 1 || Cin > 4 || KW > 8 || KH > kMaxTapsPerCta || stride < 1 || stride > 2) return false;
  const tc::StemDims d = tc::stem_dims(H, W, KH, KW, stride, pad);
  const int wstrips = (d.Q + 63) / 64;
  const int bw = ((d.Q + wstrips - 1) / wstrips + 7) / 8 * 8;
  int bh = 64 / bw;
  if (bh < 1) bh = 1;
  g->BW = bw; g->BH = bh; g->strips = (d.P + bh - 1) / bh; g->wstrips = wstrips; g->units = N * g->strips * wstrips;
  const long long stage_bytes = static_cast<long long>(mt / 32 + KH) * bw * bh * 128;
  if (stage_bytes * 2 > 200 * 1024) return false;
  g->Cin = 32; g->Cout = Cout; g->cin_tiles = 1;
  int splits = 2 * 148 / (Cout / mt);
  if (splits > g->units) splits = g->units;
  if (splits < 1) splits = 1;
  g->units_per_cta = (g->units + splits - 1) / splits;
  g->S = 1; g->RT = KH; g->pad = 0; g->stride = stride; g->stride_w = 1;
  g->layout_type = 1; g->lbo = 0; g->sbo = 512;
  g->tma_swizzle = static_cast<int>(CU_TENSOR_MAP_SWIZZLE_128B_ATOM_32B);
  return true;
}

}  // namespace

PDL_API int pdl_tc_stem_wgrad_supported(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad) {
  WgGeom g{};
  return stem_wg_geometry(N, H, W, Cin, Cout, KH, KW, stride, pad, &g) ? 1 : 0;
}

PDL_API long long pdl_tc_stem_wgrad_scratch_bytes(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad) {
  WgGeom g{};
  if (!stem_wg_geometry(N, H, W, Cin, Cout, KH, KW, stride, pad, &g)) return 0;
  return static_cast<long long>(wg_splits(g)) * KH * Cout * 32 * sizeof(float);
}

// xp: padded NHWC4 copy made by pdl_tc_stem_prepare; dy[N][P][Q][Cout] -> dw[Cout][Cin][KH][KW]
PDL_API int pdl_tc_stem_wgrad(const float* xp, const float* dy, float* dw, float* scratch, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                              int stride, int pad, int* error_flag, cudaStream_t stream) {
  WgGeom g{};
  if (!stem_wg_geometry(N, H, W, Cin, Cout, KH, KW, stride, pad, &g)) { pdl_set_error("stem wgrad: unsupported shape"); return -3; }
  const tc::StemDims d = tc::stem_dims(H, W, KH, KW, stride, pad);
  CUtensorMap ta, tb;
  if (int rc = act_map(&ta, dy, N, d.P, d.Q, Cout, g.BW, g.BH, g.tma_swizzle, 1)) return rc;
  {
    EncodeTiledFn enc = encode_tiled_wg();
    if (!enc) { pdl_set_error("cuTensorMapEncodeTiled is not available"); return -1; }
    cuuint64_t dims[4] = {32, (cuuint64_t)d.Q, (cuuint64_t)d.Hp, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)16 * stride, (cuuint64_t)d.Wp * 16, (cuuint64_t)d.Hp * d.Wp * 16};
    cuuint32_t box[4] = {32, (cuuint32_t)g.BW, (cuuint32_t)(g.BH * stride), 1};
    cuuint32_t es[4] = {1, 1, (cuuint32_t)stride, 1};
    CUresult r = enc(&tb, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(xp), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     static_cast<CUtensorMapSwizzle>(g.tma_swizzle), CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { pdl_set_error("cuTensorMapEncodeTiled(stem wgrad input) failed: %d", (int)r); return -2; }
  }
  int rc = wg_m_tile(Cout) == 128 ? launch_wg<128, 32>(ta, tb, g, 1, scratch, error_flag, stream)
                                  : launch_wg<64, 32>(ta, tb, g, 1, scratch, error_flag, stream);
  if (rc) return rc;
  const int total = Cout * Cin * KH * KW;
  stem_wgrad_finish_kernel<<<pdl_ceil_div(total, 256), 256, 0, stream>>>(scratch, wg_splits(g), Cout, Cin, KH, KW, dw);
  PDL_LAUNCH_CHECK();
  return 0;
}

// The two passes of pdl_tc_conv2d_wgrad as separate launches (stride 1), so that a caller can put the split-K GEMM and its
// deterministic second pass on different streams.  Both must see the same pdl_tc_wgrad_max_ctas setting.
PDL_API int pdl_tc_conv2d_wgrad_gemm(const float* x, const float* dy, float* scratch, int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad,
                                     int* error_flag, cudaStream_t stream) {
  return pdl_tc_conv2d_wgrad_strided(x, dy, nullptr, nullptr, nullptr, 0, scratch, N, H, W, Cin, Cout, KH, KW, 1, pad, error_flag, stream);
}
PDL_API int pdl_tc_conv2d_wgrad_finish(float* dw, float* db, const float* db_partial, int db_blocks, const float* scratch, int N, int H, int W, int Cin,
                                       int Cout, int KH, int KW, int pad, cudaStream_t stream) {
  WgGeom g{};
  if (!wg_geometry(N, H + 2 * pad - KH + 1, W + 2 * pad - KW + 1, Cin, Cout, KH, KW, pad, &g)) { pdl_set_error("tc wgrad: unsupported shape"); return -3; }
  const int total = (KH * KW * Cout * Cin + 31) / 32 * 32 + (db ? Cout * 32 : 0);
  PDL_CUDA_CHECK(launch_dependent(wgrad_finish_kernel, dim3(pdl_ceil_div(total, 256)), dim3(256), 0, stream, scratch, wg_splits(g), KH * KW, Cout, Cin, dw,
                                  db_partial, db_blocks, db));
  PDL_LAUNCH_CHECK();
  return 0;
}

// ------------------------------------------------------------------------------------------------------------------
// Second pass of a split-K weight gradient FUSED with the cross-GPU exchange of that gradient and the SGD update: the
// reference's "gradient ready -> bucket allreduce -> optimizer.step()" (launch_dist.py:91-92) for one layer in one kernel.
// Output index space (= the layer's slice of the gradient arena): [db (Cout) | dw (Cout*Cin*taps)].  CTA 0 reduces the bias
// partials, the other CTAs 256 weight outputs each (fixed split order: deterministic).  Then each CTA pushes its 256 values
// into every peer's inbox ([2 parities][world][total] floats of symmetric memory), raises ONE flag per peer for its block,
// waits for its peers' flags, sums the `world` inbox rows in rank order, and updates the parameters.  Same protocol and the
// same parity argument as the push allreduce (allreduce.cu) -- the difference is that the value pushed never exists as a
// separate tensor pass: the kernel that produces the gradient is the kernel that exchanges it.
// ------------------------------------------------------------------------------------------------------------------
namespace {
struct FinishComm {
  float* inbox[pdl_comm::kMaxWorld];
  uint32_t* pad[pdl_comm::kMaxWorld];
  pdl_comm::CommState* state;
  float* raw;    // local gradient out (arena `in` slice), [db | dw]
  float* avg;    // averaged gradient out
  float* param;  // parameters / momentum in the same order (nullptr: average only)
  float* mom;
  SgdHyper* hyper;
  int rank, world, channel;
  float scale;
  unsigned long long timeout_ns;
};

__global__ void __launch_bounds__(256) wgrad_finish_comm_kernel(const float* __restrict__ partial, int splits, int taps, int Cout, int Cin,
                                                                const float* __restrict__ db_partial, int db_blocks, FinishComm fc) {
  using namespace pdl_comm;
  launch_dependents();
  grid_dependency_wait();
  const int per_split = taps * Cout * Cin;
  const int total = Cout + per_split;
  // ---- second pass: this thread's output ----
  int i = -1;          // index in [db | dw]
  float mine = 0.f;
  if (blockIdx.x == 0) {  // bias gradient: 256 / Cout... threads per channel stride over the partial rows, fixed shuffle tree
    const int tpc = 256 / Cout > 0 ? 256 / Cout : 1;  // Cout <= 256
    const int co = threadIdx.x / tpc, sub = threadIdx.x % tpc;
    float acc = 0.f;
    if (co < Cout)
      for (int b = sub; b < db_blocks; b += tpc) acc += db_partial[b * Cout + co];
    for (int o = tpc >> 1; o; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
    if (co < Cout && sub == 0) { i = co; mine = acc; }
  } else {
    const int e = (blockIdx.x - 1) * 256 + threadIdx.x;  // output layout [co][ci][tap]: coalesced writes
    if (e < per_split) {
      const int tap = e % taps, ci = (e / taps) % Cin, co = e / (taps * Cin);
      const float* src = partial + (static_cast<long long>(tap) * Cout + co) * Cin + ci;
      float acc = 0.f;
#pragma unroll 8
      for (int s = 0; s < splits; ++s) acc += __ldg(src + static_cast<long long>(s) * per_split);
      i = Cout + e;
      mine = acc;
    }
  }
  if (i >= 0) fc.raw[i] = mine;
  // ---- exchange ----
  const int W = fc.world, r = fc.rank, blk = blockIdx.x;
  float sum = mine;
  if (W > 1) {
    if (__syncthreads_or(fc.state->error != 0)) return;  // an earlier step lost a peer: leave parameters untouched
    const uint32_t ep = fc.state->epoch[fc.channel];
    const size_t row = (static_cast<size_t>(ep & 1u) * W + r) * total;
    if (i >= 0) {
#pragma unroll
      for (int q = 0; q < kMaxWorld; ++q)
        if (q < W && q != r) asm volatile("st.global.relaxed.sys.f32 [%0], %1;" ::"l"(fc.inbox[q] + row + i), "f"(mine) : "memory");
    }
    __syncthreads();
    int failed = 0;
    if (threadIdx.x < W && threadIdx.x != r) {
      st_release_sys(fc.pad[threadIdx.x] + flag_index(fc.channel, blk, r), ep + 1);
      failed = poll_flag(fc.pad[r] + flag_index(fc.channel, blk, threadIdx.x), ep + 1, fc.timeout_ns, fc.state) ? 0 : 1;
    }
    if (__syncthreads_or(failed)) return;
    if (i >= 0) {
      const float* base = fc.inbox[r] + static_cast<size_t>(ep & 1u) * W * total + i;
      float v[kMaxWorld];
#pragma unroll
      for (int q = 0; q < kMaxWorld; ++q)
        if (q < W && q != r) asm volatile("ld.global.relaxed.sys.f32 %0, [%1];" : "=f"(v[q]) : "l"(base + static_cast<size_t>(q) * total) : "memory");
      sum = r == 0 ? mine : v[0];
#pragma unroll
      for (int q = 1; q < kMaxWorld; ++q)  // fixed rank order: every rank computes the same bits
        if (q < W) sum += (q == r ? mine : v[q]);
    }
  }
  if (i >= 0) {
    const float avg = sum * fc.scale;
    fc.avg[i] = avg;
    if (fc.param != nullptr) {
      const SgdHyper h = *fc.hyper;
      const bool has_mom = fc.mom != nullptr && h.momentum != 0.0f;
      float m = has_mom ? fc.mom[i] : 0.f;
      fc.param[i] = sgd_update(fc.param[i], avg, &m, h, has_mom);
      if (has_mom) fc.mom[i] = m;
    }
  }
  // the last CTA to finish advances the epoch (all CTAs have read it by then) and clears the first-step flag
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    const uint32_t done = atomicAdd(&fc.state->finished[fc.channel], 1u);
    if (done == gridDim.x - 1) {
      fc.state->finished[fc.channel] = 0;
      if (W > 1) fc.state->epoch[fc.channel] += 1;
      if (fc.param != nullptr) fc.hyper->first_step = 0;
      __threadfence();
    }
  }
}
}  // namespace

// Floats of symmetric memory per rank the exchange needs: [2][world][Cout + Cout*Cin*KH*KW].
PDL_API long long pdl_wgrad_finish_comm_inbox_floats(int world, int Cin, int Cout, int KH, int KW) {
  return 2ll * world * (static_cast<long long>(Cout) + static_cast<long long>(Cout) * Cin * KH * KW);
}
PDL_API int pdl_wgrad_finish_comm_blocks(int Cin, int Cout, int KH, int KW) { return 1 + pdl_ceil_div(static_cast<long long>(Cout) * Cin * KH * KW, 256); }

// Second pass of pdl_tc_conv2d_wgrad_gemm + exchange + SGD.  raw/avg/param/mom point at the layer's [db | dw] slice (bias first,
// contiguous); inbox/pad: `world` peer-mapped pointers (pad: a signal pad; `channel` of it is used with one flag block per
// CTA, so pdl_wgrad_finish_comm_blocks() must not exceed pdl_comm_max_blocks()); state: communicator state whose `channel` is
// used by nothing else in flight.  world == 1: no exchange.  Must see the same pdl_tc_wgrad_max_ctas setting as the GEMM pass.
PDL_API int pdl_tc_conv2d_wgrad_finish_comm(const float* scratch, const float* db_partial, int db_blocks, int N, int H, int W, int Cin, int Cout,
                                            int KH, int KW, int pad_, float* raw, float* avg, float* param, float* mom, void* hyper, int rank,
                                            int world, int channel, float* const* inbox, uint32_t* const* pad, void* state, double scale,
                                            double timeout_s, cudaStream_t stream) {
  WgGeom g{};
  if (!wg_geometry(N, H + 2 * pad_ - KH + 1, W + 2 * pad_ - KW + 1, Cin, Cout, KH, KW, pad_, &g)) { pdl_set_error("tc wgrad: unsupported shape"); return -3; }
  if (Cout > 256 || (256 % Cout) != 0) { pdl_set_error("wgrad finish+exchange: Cout must divide 256"); return -3; }
  const int blocks = pdl_wgrad_finish_comm_blocks(Cin, Cout, KH, KW);
  if (world < 1 || world > pdl_comm::kMaxWorld || channel < 0 || channel >= pdl_comm::kMaxChannels || !state) {
    pdl_set_error("wgrad finish+exchange: bad world / channel / state");
    return -1;
  }
  if (world > 1 && (blocks > pdl_comm::kMaxBlocks || !inbox || !pad)) { pdl_set_error("wgrad finish+exchange: too many blocks or peer pointers missing"); return -1; }
  FinishComm fc{};
  for (int q = 0; q < world && world > 1; ++q) { fc.inbox[q] = inbox[q]; fc.pad[q] = pad[q]; }
  fc.state = static_cast<pdl_comm::CommState*>(state);
  fc.raw = raw; fc.avg = avg; fc.param = param; fc.mom = mom; fc.hyper = static_cast<SgdHyper*>(hyper);
  fc.rank = rank; fc.world = world; fc.channel = channel; fc.scale = static_cast<float>(scale);
  fc.timeout_ns = static_cast<unsigned long long>(timeout_s * 1e9);
  PDL_CUDA_CHECK(launch_dependent(wgrad_finish_comm_kernel, dim3(blocks), dim3(256), 0, stream, scratch, wg_splits(g), KH * KW, Cout, Cin, db_partial,
                                  db_blocks, fc));
  PDL_LAUNCH_CHECK();
  return 0;
}
