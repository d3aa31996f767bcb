// Implicit-GEMM convolution on the 5th-gen tensor cores: TMA-fed, tcgen05.mma (kind::tf32) with the
// accumulator in TMEM, warp-specialised (TMA producer / single-thread MMA issuer / epilogue warps).
//
// Replaces the cuDNN TF32 convolutions the reference's ConvNet runs (launch_dist.py:15-25 forward,
// backward through loss.backward() at :91):
//   forward  y[m, co]  = sum_{r,s,ci} x[pix(m)+(r,s), ci] * w[co, ci, r, s]   (+bias, ReLU, 2x2 max-pool fused)
//   dgrad    dx[m, ci] = sum_{r,s,co} dy[pix(m)-(r,s), co] * w[co, ci, r, s]   (TMA zero-fills the halo)
// Activations are NHWC.  One M tile = a 4-D TMA box (32 channels x BW x BH x BN pixels <= 128 rows)
// landing as 128-byte rows in a SWIZZLE_128B K-major operand; the filter tap only shifts the box
// origin, so no im2col buffer ever exists.  Weights are pre-packed per tap as [tap][n][k] (K-major).
// fp32 storage everywhere, TF32 multiplies, fp32 accumulation (the reference's convs are TF32 too).
#include "common.cuh"
#include "tc_common.cuh"

#include <cstdlib>

namespace {

constexpr int kThreads = 384;   // 4 control warps (TMA, MMA, TMEM alloc, spare) + 8 epilogue warps
constexpr int kEpiWarps = 8;
constexpr int A_BYTES = 128 * 128;  // 128 rows x 32 tf32

enum Epi : int { EPI_STORE = 0, EPI_RELU_POOL = 1, EPI_POOL_BWD = 2 };

struct TcGeom {
  int num_tiles;
  int BW, BH, BN;        // TMA box extents in pixels; rows = BW*BH*BN <= 128
  int tiles_per_img;     // row strips per image when BN == 1
  int R, S, kchunks;     // filter taps, 32-wide K chunks per tap
  int a_w0, a_h0;        // input coordinate read by output pixel (0,0) at tap (0,0)
  int tap_sign;          // +1: coord = base + tap (forward); -1: coord = base - tap (dgrad)
  int Nimg, Hout, Wout;  // output extent (Wout == BW)
  int in_stride;         // convolution stride (forward): box origin row = h0 * in_stride + a_h0 + tap
  int tap_r0, tap_s0, tap_step, S_full;  // taps visited: (tap_r0 + r*tap_step, tap_s0 + s*tap_step) of an S_full-wide filter
  int out_stride, out_oh, out_ow, Hc;    // output pixel = (tile row * out_stride + out_oh, col * out_stride + out_ow); Hc tile-grid rows
  int n_tiles, ldc;      // N_GEMM-wide column tiles; ldc = total output channels (tile index = m_tile * n_tiles + n_tile)
  const float* w_base;   // the packed weights behind tmB ([w_rows][w_K]): lets the launcher re-encode the map for CTA pairs
  int w_rows, w_K;
};

struct EpiArgs {
  float* out;
  uint8_t* idx;
  const float* bias;
  int relu;
  int pool_stride;
  int tma_store;     // EPI_STORE: hand the tile to TMA through the swizzled staging buffer (dense outputs only)
  int* error;
  long long* trace;  // optional: clock64() of CTA 0 at pipeline milestones (profiling aid)
  const float* pooled;  // EPI_POOL_BWD: the pooled activations (ReLU mask) ...
  float* dbpart;        // ... and where the per-tile column sums go; `idx` is then the pool's arg-max, `out` is dz
  float* bn_partial;    // EPI_STORE + tma_store: per-CTA (count, mean, M2) of every output channel, [grid / n_tiles][ldc][3]
  const float* add_src; // EPI_STORE + tma_store: tensor of the output's shape added to the result (skip-connection gradient)
};

template <int N_GEMM>
struct Smem {
  static constexpr int B_BYTES = N_GEMM * 128;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  // what fits next to the staging tile.  N = 256 (48 KB stages) exists because these kernels are bound by aggregate L2
  // traffic: one A tile then serves twice the columns (24 KB per 128x128x32 of MACs instead of 32 KB).
  static constexpr int STAGES = N_GEMM == 256 ? 3 : N_GEMM == 128 ? 4 : (N_GEMM == 64 ? 6 : 8);
  static constexpr int PASS_N = N_GEMM > 128 ? 128 : N_GEMM;  // columns the epilogue stages (and stores) at a time
  static constexpr int STAGING_LD = PASS_N + 4;
  static constexpr int STAGING_BYTES = 128 * STAGING_LD * 4;
  static constexpr int BAR_OFFSET = STAGES * STAGE_BYTES + STAGING_BYTES;
  static constexpr int TOTAL = BAR_OFFSET + 256 + 512 + 512 + 1024;  // + barriers + bias row + pool table + alignment slack
};

// CL = 2: the kernel runs as clusters of two CTAs that work on two M tiles of the SAME column block in lock step and share
// the weight tile: each CTA fetches half of it and TMA-multicasts that half into both CTAs' stage, so every weight byte
// crosses L2 -> SM once per pair instead of once per CTA (these kernels are bound by exactly that traffic).  A stage may be
// refilled only when BOTH CTAs' MMAs have retired, hence the multicast commit on the `empty` barriers (count 2).
// Tiles are paired as (v, v ^ 1): virtual tile v -> pair p = v >> 1 -> (m tile 2 * (p / n_tiles) + (v & 1), n tile p % n_tiles).
template <int CL>
__device__ __forceinline__ void decode_tile(int tile, int n_tiles, int& mt, int& nt) {
  if (CL == 1) { mt = tile / n_tiles; nt = tile % n_tiles; }
  else { const int p = tile >> 1; mt = 2 * (p / n_tiles) + (tile & 1); nt = p % n_tiles; }
}

template <int N_GEMM, int EPI, int CL>
__global__ void __launch_bounds__(kThreads, 1)
tc_conv_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB, const __grid_constant__ CUtensorMap tmOut,
               const TcGeom g, const EpiArgs e) {
  using SM = Smem<N_GEMM>;
  constexpr int kStages = SM::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
  float* staging = reinterpret_cast<float*>(smem + kStages * SM::STAGE_BYTES);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + SM::BAR_OFFSET);
  uint64_t* empty = full + kStages;
  uint64_t* tfull = empty + kStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  volatile int* failed = reinterpret_cast<volatile int*>(tmem_slot + 1);
  float* sbias = reinterpret_cast<float*>(smem + SM::BAR_OFFSET + 256);       // N_GEMM floats
  int* pool_row = reinterpret_cast<int*>(smem + SM::BAR_OFFSET + 256 + 512);   // <= 128 pooled pixels -> staging row

  const int warp = threadIdx.x / 32, lane = threadIdx.x % 32;
#define PDL_TRACE(slot) do { if (e.trace && blockIdx.x == 0) e.trace[slot] = clock64(); } while (0)
  if (threadIdx.x == 0) PDL_TRACE(0);
  constexpr uint32_t kTmemCols = (2 * N_GEMM <= 32) ? 32 : (2 * N_GEMM <= 64) ? 64 : (2 * N_GEMM <= 128) ? 128 : (2 * N_GEMM <= 256) ? 256 : 512;

  if (threadIdx.x == 0) {
    tc::prefetch_tmap(&tmA);
    tc::prefetch_tmap(&tmB);
    if (EPI == EPI_STORE && e.tma_store) tc::prefetch_tmap(&tmOut);
    for (int i = 0; i < kStages; ++i) { tc::mbar_init(&full[i], 1); tc::mbar_init(&empty[i], CL); }
    for (int i = 0; i < 2; ++i) { tc::mbar_init(&tfull[i], 1); tc::mbar_init(&tempty[i], kEpiWarps); }
    *failed = 0;
    tc::fence_barrier_init();
  }
  if (warp == 2) tc::tmem_alloc(tmem_slot, kTmemCols);
  // Everything above ran (possibly) while the previous kernel of the stream was still draining; from here on we read
  // what it produced.  Our own successor may start its prologue as soon as all of our CTAs are resident.
  launch_dependents();
  grid_dependency_wait();
  // per-CTA constants of the epilogue, computed once: bias row, pooled pixel -> first staging row of its window
  const int ps = e.pool_stride > 0 ? e.pool_stride : 1;
  const int PH = (g.BH - 2) / ps + 1, PW = (g.BW - 2) / ps + 1;
  if (threadIdx.x >= 128) {
    const int t = threadIdx.x - 128;
    if (EPI == EPI_RELU_POOL && t < N_GEMM) sbias[t] = e.bias ? e.bias[t] : 0.f;
    if (EPI == EPI_RELU_POOL && t < g.BN * PH * PW) {
      const int img = t / (PH * PW), pp = t % (PH * PW);
      pool_row[t] = (img * g.BH + (pp / PW) * ps) * g.BW + (pp % PW) * ps;
    }
  }
  tc::tc_fence_before();
  __syncthreads();
  if (CL > 1) tc::cluster_sync();  // the peer's barriers exist before anything is multicast at them
  tc::tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t crank = CL > 1 ? tc::cluster_ctarank() : 0;
  if (threadIdx.x == 0) PDL_TRACE(1);

  const int ksteps = g.R * g.S * g.kchunks;
  const int rows = g.BW * g.BH * g.BN;
  const uint32_t stage_tx = static_cast<uint32_t>(rows * 128 + SM::B_BYTES);

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (tc::elect_one()) {
      int stage = 0;
      uint32_t phase = 0;
      bool ok = true;
      for (int tile = blockIdx.x; tile < g.num_tiles && ok; tile += gridDim.x) {
        int mt, nt;
        decode_tile<CL>(tile, g.n_tiles, mt, nt);
        const int n0 = g.BN > 1 ? mt * g.BN : mt / g.tiles_per_img;
        const int h0 = g.BN > 1 ? 0 : (mt % g.tiles_per_img) * g.BH;
        int tap = 0;
        for (int r = 0; r < g.R && ok; ++r)
          for (int s = 0; s < g.S && ok; ++s, ++tap)
            for (int kc = 0; kc < g.kchunks; ++kc) {
              if (!tc::mbar_wait(&empty[stage], phase ^ 1)) { *failed = 1; ok = false; break; }
              uint8_t* a_dst = smem + stage * SM::STAGE_BYTES;
              tc::mbar_expect_tx(&full[stage], stage_tx);
              tc::tma_load_4d(a_dst, &tmA, &full[stage], kc * 32, g.a_w0 + g.tap_sign * s, h0 * g.in_stride + g.a_h0 + g.tap_sign * r, n0);
              const int wtap = (g.tap_r0 + r * g.tap_step) * g.S_full + g.tap_s0 + s * g.tap_step;  // row block of the packed weights
              if (CL == 1)
                tc::tma_load_2d(a_dst + A_BYTES, &tmB, &full[stage], kc * 32, wtap * g.ldc + nt * N_GEMM);
              else  // my half of the weight tile, into both CTAs' stage (tmB's box is N_GEMM / 2 rows here)
                tc::tma_load_2d_multicast(a_dst + A_BYTES + crank * (SM::B_BYTES / 2), &tmB, &full[stage], kc * 32,
                                          wtap * g.ldc + nt * N_GEMM + static_cast<int>(crank) * (N_GEMM / 2), static_cast<uint16_t>(0x3));
              if (tile == blockIdx.x && tap == 0 && kc == 0) PDL_TRACE(2);
              if (++stage == kStages) { stage = 0; phase ^= 1; }
            }
      }
      PDL_TRACE(3);
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    constexpr uint32_t idesc = tc::idesc_tf32(128, N_GEMM, 0, 0);
    int stage = 0, acc = 0;
    uint32_t phase = 0, acc_phase = 0;
    for (int tile = blockIdx.x; tile < g.num_tiles && !*failed; tile += gridDim.x) {
      if (!tc::mbar_wait(&tempty[acc], acc_phase ^ 1)) { *failed = 1; break; }
      tc::tc_fence_after();
      const uint32_t d_tmem = tmem_base + acc * N_GEMM;
      for (int ks = 0; ks < ksteps; ++ks) {
        if (!tc::mbar_wait(&full[stage], phase)) { *failed = 1; break; }
        tc::tc_fence_after();
        if (tile == blockIdx.x && ks == 0 && lane == 0) PDL_TRACE(4);
        if (tile == blockIdx.x && ks == ksteps - 1 && lane == 0) PDL_TRACE(5);
        if (tc::elect_one()) {
          const uint32_t a_addr = tc::smem_u32(smem + stage * SM::STAGE_BYTES);
          const uint32_t b_addr = a_addr + A_BYTES;
#pragma unroll
          for (int k = 0; k < 4; ++k) {  // 4 x K=8 tf32 = one 128-byte swizzle row
            tc::mma_tf32(d_tmem, tc::smem_desc_sw128(a_addr + k * 32, 0, 1024), tc::smem_desc_sw128(b_addr + k * 32, 0, 1024), idesc,
                         (ks | k) != 0);
          }
          if (CL == 1) tc::mma_commit(&empty[stage]);         // frees the smem slot when these MMAs retire
          else tc::mma_commit_multicast(&empty[stage], static_cast<uint16_t>(0x3));  // ... in both CTAs: the slot is filled by both
          if (ks == ksteps - 1) tc::mma_commit(&tfull[acc]);  // accumulator complete -> epilogue
        }
        __syncwarp();
        if (++stage == kStages) { stage = 0; phase ^= 1; }
      }
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
  } else if (warp >= 4) {
    // ===================== epilogue: 8 warps; warp pair (q, q+4) shares TMEM lane quarter q and splits the columns =====================
    const int et = threadIdx.x - 128;           // 0..255
    const int ew = warp - 4, quarter = ew & 3, half = ew >> 2;
    const int row = quarter * 32 + lane;        // accumulator row = output pixel of the tile
    constexpr int PN = SM::PASS_N, NPASS = N_GEMM / PN;  // the epilogue handles PN columns at a time (two passes for N = 256)
    constexpr int HALF = PN / 2;                // columns per warp (per pass)
    constexpr int CH = HALF < 32 ? HALF : 32;   // columns per TMEM load
    int acc = 0;
    uint32_t acc_phase = 0;
    float bn_sum[NPASS], bn_sq[NPASS];  // fused BatchNorm statistics (see e.bn_partial)
#pragma unroll
    for (int p = 0; p < NPASS; ++p) { bn_sum[p] = 0.f; bn_sq[p] = 0.f; }
    int bn_count = 0;
    for (int tile = blockIdx.x; tile < g.num_tiles; tile += gridDim.x) {
      if (*failed || !tc::mbar_wait(&tfull[acc], acc_phase)) { *failed = 1; break; }
      tc::tc_fence_after();
      if (tile == blockIdx.x && et == 0) PDL_TRACE(6);
      int mt, nt;
      decode_tile<CL>(tile, g.n_tiles, mt, nt);
      const int n0 = g.BN > 1 ? mt * g.BN : mt / g.tiles_per_img;
      const int h0 = g.BN > 1 ? 0 : (mt % g.tiles_per_img) * g.BH;
      const uint32_t t_row = tmem_base + acc * N_GEMM + half * HALF + (static_cast<uint32_t>(quarter * 32) << 16);
      if (EPI == EPI_STORE && e.tma_store) {
       // registers -> swizzled staging tile (one [128 rows][128 B] block per 32 channels) -> one TMA store per block.
       // The store clips rows / images past the tensor edge, so no per-row predicate is needed.
#pragma unroll
       for (int pass = 0; pass < NPASS; ++pass) {
        const int pc0 = nt * N_GEMM + pass * PN;  // first output channel of this pass
        const uint32_t t_row = tmem_base + acc * N_GEMM + pass * PN + half * HALF + (static_cast<uint32_t>(quarter * 32) << 16);
        if (et == 0) tc::bulk_wait_read_all();  // the previous stores have drained the staging buffer
        asm volatile("bar.sync 1, 256;" ::: "memory");
        const float* bias_t = e.bias ? e.bias + pc0 + half * HALF : nullptr;
        // optional second addend with the output's layout (the skip connection's gradient, when this is the dgrad of the
        // first convolution of a residual block): out = acc + add_src.  Rows past the tensor edge are clipped by the store
        // and must not be read here.
        const float* add_row = nullptr;
        if (e.add_src) {
          const int bw_ = row % g.BW, t2_ = row / g.BW, bh_ = t2_ % g.BH, bn_ = t2_ / g.BH;
          if (row < rows && (n0 + bn_) < g.Nimg && (h0 + bh_) < g.Hc)
            add_row = e.add_src + ((static_cast<long long>(n0 + bn_) * g.Hout + h0 + bh_) * g.Wout + bw_) * g.ldc + pc0 + half * HALF;
        }
#pragma unroll 1
        for (int c = 0; c < HALF; c += CH) {
          float v[CH];
          tc::tmem_ld<CH>(t_row + c, v);
          const int col0 = half * HALF + c;
          uint8_t* dst_row = reinterpret_cast<uint8_t*>(staging) + (col0 / 32) * A_BYTES + row * 128;
          const int j0 = (col0 % 32) / 4;
#pragma unroll
          for (int j = 0; j < CH; j += 4) {
            const float4 bb = bias_t ? *reinterpret_cast<const float4*>(bias_t + c + j) : make_float4(0.f, 0.f, 0.f, 0.f);
            float4 o = make_float4(v[j] + bb.x, v[j + 1] + bb.y, v[j + 2] + bb.z, v[j + 3] + bb.w);
            if (add_row) {
              const float4 aa = __ldg(reinterpret_cast<const float4*>(add_row + c + j));
              o.x += aa.x; o.y += aa.y; o.z += aa.z; o.w += aa.w;
            }
            if (e.relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
            *reinterpret_cast<float4*>(dst_row + (((j0 + j / 4) ^ (row & 7)) << 4)) = o;
          }
        }
        if (pass == NPASS - 1) {
          tc::tc_fence_before();
          __syncwarp();
          if (lane == 0) tc::mbar_arrive(&tempty[acc]);  // accumulator drained: the next tile's MMAs may start
        }
        tc::fence_proxy_async_smem();
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (et == 0) {
#pragma unroll
          for (int cc = 0; cc < PN / 32; ++cc)
            tc::tma_store_4d(&tmOut, reinterpret_cast<uint8_t*>(staging) + cc * A_BYTES, pc0 + cc * 32, 0, h0, n0);
          tc::bulk_commit();
        }
        if (e.bn_partial) {
          // BatchNorm statistics of this output, straight from the staged tile (saves the normalisation layer a full
          // read of the tensor).  The grid is a multiple of n_tiles, so every tile of this CTA has the same column
          // block and the sums live in two registers per thread: thread = (column, row slice).  Only the rows that exist
          // in the output count: a ragged last strip / a last tile with fewer images still computes the rows past the
          // tensor edge (from the halo of valid input), the TMA store merely clips them.  Rows are ordered image, row,
          // column, so the valid ones are a prefix of the tile.
          constexpr int PARTS = 256 / PN;
          const int col = et % PN, part = et / PN;
          const uint8_t* src = reinterpret_cast<const uint8_t*>(staging) + (col / 32) * A_BYTES + (col % 4) * 4;
          const int piece = (col % 32) / 4;
          const int vh = min(g.BH, g.Hc - h0), vn = min(g.BN, g.Nimg - n0);
          const int valid_rows = g.BW * (vh > 0 ? vh : 0) * (vn > 0 ? vn : 0);
          float s_ = 0.f, q_ = 0.f;
#pragma unroll 4
          for (int r = part; r < valid_rows; r += PARTS) {
            const float v = *reinterpret_cast<const float*>(src + r * 128 + ((piece ^ (r & 7)) << 4));
            s_ += v;
            q_ = fmaf(v, v, q_);
          }
          bn_sum[pass] += s_;
          bn_sq[pass] += q_;
          if (pass == 0) bn_count += valid_rows;
        }
       }  // pass
      } else if (EPI == EPI_STORE) {
        const int bw = row % g.BW, t2 = row / g.BW, bh = t2 % g.BH, bn = t2 / g.BH;
        // output pixel of this row: the tile grid (Hc x BW) is the full image, or one parity class of it (strided dgrad)
        const int oh = (h0 + bh) * g.out_stride + g.out_oh, ow = bw * g.out_stride + g.out_ow;
        const bool valid = row < rows && (n0 + bn) < g.Nimg && (h0 + bh) < g.Hc && oh < g.Hout && ow < g.Wout;
        float* dst = e.out + ((static_cast<long long>(n0 + bn) * g.Hout + oh) * g.Wout + ow) * g.ldc + nt * N_GEMM + half * HALF;
        const float* bias_t = e.bias ? e.bias + nt * N_GEMM + half * HALF : nullptr;
#pragma unroll 1
        for (int c = 0; c < HALF; c += CH) {
          float v[CH];
          tc::tmem_ld<CH>(t_row + c, v);
          if (valid) {
#pragma unroll
            for (int j = 0; j < CH; j += 4) {
              const float4 bb = bias_t ? *reinterpret_cast<const float4*>(bias_t + c + j) : make_float4(0.f, 0.f, 0.f, 0.f);
              float4 o = make_float4(v[j] + bb.x, v[j + 1] + bb.y, v[j + 2] + bb.z, v[j + 3] + bb.w);
              if (e.relu) { o.x = fmaxf(o.x, 0.f); o.y = fmaxf(o.y, 0.f); o.z = fmaxf(o.z, 0.f); o.w = fmaxf(o.w, 0.f); }
              *reinterpret_cast<float4*>(dst + c + j) = o;
            }
          }
        }
        tc::tc_fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&tempty[acc]);
      } else if (EPI == EPI_POOL_BWD) {
        // The tile is the gradient of a 2x2-max-pooled tensor (whole images).  Stage it, then route it back through the
        // arg-max of each window with the ReLU mask folded in (windows whose pooled value is <= 0 pass nothing):
        // dz[n][ZH][ZW][ldc], rounded to TF32 for the next tensor-core op, plus the tile's column sums (the bias-gradient
        // partials of the layer below).  Saves a kernel on the critical path of the backward pass.
#pragma unroll 1
        for (int c = 0; c < HALF; c += CH) {
          float v[CH];
          tc::tmem_ld<CH>(t_row + c, v);
#pragma unroll
          for (int j = 0; j < CH; j += 4)
            *reinterpret_cast<float4*>(staging + row * SM::STAGING_LD + half * HALF + c + j) = make_float4(v[j], v[j + 1], v[j + 2], v[j + 3]);
        }
        tc::tc_fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&tempty[acc]);
        asm volatile("bar.sync 1, 256;" ::: "memory");
        constexpr int C4 = N_GEMM / 4, LANES = 256 / C4;
        const int c4 = et % C4, pl = et / C4;
        const int ZH = (g.BH - 1) * ps + 2, ZW = (g.BW - 1) * ps + 2, zpix = ZH * ZW;
        const int nimg = min(g.BN, g.Nimg - n0);
        const int ch0 = nt * N_GEMM + c4 * 4;
        float4 colsum = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 2
        for (int px = pl; px < nimg * zpix; px += LANES) {
          const int img = px / zpix, rem = px % zpix, h = rem / ZW, w = rem % ZW;
          float4 a4 = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int k = 0; k < 4; ++k) {  // the window whose element (k/2, k%2) is this pixel
            const int ph_ = h - (k >> 1), pw_ = w - (k & 1);
            const bool valid = ph_ >= 0 && pw_ >= 0 && (ph_ % ps) == 0 && (pw_ % ps) == 0 && (ph_ / ps) < g.BH && (pw_ / ps) < g.BW;
            const int prow = valid ? (img * g.BH + ph_ / ps) * g.BW + pw_ / ps : 0;  // branch-free: the loads overlap
            const float4 gv = *reinterpret_cast<const float4*>(staging + prow * SM::STAGING_LD + c4 * 4);
            const long long o = (static_cast<long long>(n0) * g.BH * g.BW + prow) * g.ldc + ch0;
            const float4 pv = __ldg(reinterpret_cast<const float4*>(e.pooled + o));
            const uchar4 am = __ldg(reinterpret_cast<const uchar4*>(e.idx + o));
            a4.x += (valid && am.x == k && pv.x > 0.f) ? gv.x : 0.f;
            a4.y += (valid && am.y == k && pv.y > 0.f) ? gv.y : 0.f;
            a4.z += (valid && am.z == k && pv.z > 0.f) ? gv.z : 0.f;
            a4.w += (valid && am.w == k && pv.w > 0.f) ? gv.w : 0.f;
          }
          colsum.x += a4.x; colsum.y += a4.y; colsum.z += a4.z; colsum.w += a4.w;
          *reinterpret_cast<float4*>(e.out + ((static_cast<long long>(n0 + img) * ZH + h) * ZW + w) * g.ldc + ch0) =
              make_float4(round_tf32(a4.x), round_tf32(a4.y), round_tf32(a4.z), round_tf32(a4.w));
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");  // everyone is done reading the staged gradients
        *reinterpret_cast<float4*>(staging + pl * SM::STAGING_LD + c4 * 4) = colsum;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (et < N_GEMM) {
          float s = 0.f;
#pragma unroll
          for (int i = 0; i < LANES; ++i) s += staging[i * SM::STAGING_LD + et];  // fixed order
          e.dbpart[static_cast<long long>(mt) * g.ldc + nt * N_GEMM + et] = s;
        }
        asm volatile("bar.sync 1, 256;" ::: "memory");  // staging may be overwritten by the next tile
      } else {
        // bias + ReLU into the staging tile, then 2x2 max-pool (stride 1 or 2) with first-max argmax
#pragma unroll 1
        for (int c = 0; c < HALF; c += CH) {
          float v[CH];
          tc::tmem_ld<CH>(t_row + c, v);
          if (tile == blockIdx.x && et == 0 && c == 0) PDL_TRACE(9);
#pragma unroll
          for (int j = 0; j < CH; j += 4) {
            const float4 bb = *reinterpret_cast<const float4*>(sbias + half * HALF + c + j);
            float4 o = make_float4(fmaxf(v[j] + bb.x, 0.f), fmaxf(v[j + 1] + bb.y, 0.f), fmaxf(v[j + 2] + bb.z, 0.f), fmaxf(v[j + 3] + bb.w, 0.f));
            *reinterpret_cast<float4*>(staging + row * SM::STAGING_LD + half * HALF + c + j) = o;
          }
        }
        if (tile == blockIdx.x && et == 0) PDL_TRACE(10);
        tc::tc_fence_before();
        __syncwarp();
        if (lane == 0) tc::mbar_arrive(&tempty[acc]);  // TMEM buffer is free: the next tile's MMAs may start
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (tile == blockIdx.x && et == 0) PDL_TRACE(11);
        constexpr int C4 = N_GEMM / 4;
        const int c4 = et % C4;
        const int npp = min(g.BN, g.Nimg - n0) * PH * PW;  // pooled pixels of this tile that belong to real images
        float* out_base = e.out + static_cast<long long>(n0) * PH * PW * N_GEMM + c4 * 4;
        uint8_t* idx_base = e.idx + static_cast<long long>(n0) * PH * PW * N_GEMM + c4 * 4;
#pragma unroll 2
        for (int pp = et / C4; pp < npp; pp += 256 / C4) {
          const float* base = staging + pool_row[pp] * SM::STAGING_LD + c4 * 4;
          const float4 v00 = *reinterpret_cast<const float4*>(base);
          const float4 v01 = *reinterpret_cast<const float4*>(base + SM::STAGING_LD);
          const float4 v10 = *reinterpret_cast<const float4*>(base + g.BW * SM::STAGING_LD);
          const float4 v11 = *reinterpret_cast<const float4*>(base + (g.BW + 1) * SM::STAGING_LD);
          float4 best = v00;
          uchar4 arg = make_uchar4(0, 0, 0, 0);
#define PDL_POOL_STEP(V, K)                                   \
  if (V.x > best.x) { best.x = V.x; arg.x = K; }              \
  if (V.y > best.y) { best.y = V.y; arg.y = K; }              \
  if (V.z > best.z) { best.z = V.z; arg.z = K; }              \
  if (V.w > best.w) { best.w = V.w; arg.w = K; }
          PDL_POOL_STEP(v01, 1)
          PDL_POOL_STEP(v10, 2)
          PDL_POOL_STEP(v11, 3)
#undef PDL_POOL_STEP
          best = make_float4(round_tf32(best.x), round_tf32(best.y), round_tf32(best.z), round_tf32(best.w));  // next layer's MMA operand
          *reinterpret_cast<float4*>(out_base + static_cast<long long>(pp) * N_GEMM) = best;
          *reinterpret_cast<uchar4*>(idx_base + static_cast<long long>(pp) * N_GEMM) = arg;
        }
        if (tile == blockIdx.x && et == 0) PDL_TRACE(12);
        asm volatile("bar.sync 1, 256;" ::: "memory");  // staging may be overwritten by the next tile
      }
      if (tile == blockIdx.x && et == 0) PDL_TRACE(7);
      if (++acc == 2) { acc = 0; acc_phase ^= 1; }
    }
    if (EPI == EPI_STORE && e.tma_store && et == 0) tc::bulk_wait_read_all();  // smem must outlive the last store's reads
    if (EPI == EPI_STORE && e.tma_store && e.bn_partial) {
      // this CTA's (count, mean, M2) per column -> bn_partial[split][C][3], split = blockIdx.x / n_tiles
      constexpr int PARTS = 256 / PN;
      const int col = et % PN, part = et / PN;
#pragma unroll
      for (int pass = 0; pass < NPASS; ++pass) {
        asm volatile("bar.sync 1, 256;" ::: "memory");  // all stores drained (or the previous pass read): staging is free
        staging[(part * 2 + 0) * PN + col] = bn_sum[pass];
        staging[(part * 2 + 1) * PN + col] = bn_sq[pass];
        asm volatile("bar.sync 1, 256;" ::: "memory");
        if (et < PN) {
          float s = 0.f, q = 0.f;
#pragma unroll
          for (int p = 0; p < PARTS; ++p) { s += staging[(p * 2) * PN + et]; q += staging[(p * 2 + 1) * PN + et]; }  // fixed order
          const float n = static_cast<float>(bn_count);
          const float mean = n > 0.f ? s / n : 0.f;
          // every tile of this CTA has the same column block (the grid is a multiple of CL * n_tiles): its index, and this
          // CTA's row among the grid / n_tiles CTAs that share it
          const int cta = CL == 1 ? blockIdx.x : blockIdx.x >> 1;
          const int nt_mine = cta % g.n_tiles;
          const int split = CL == 1 ? cta / g.n_tiles : (cta / g.n_tiles) * 2 + (blockIdx.x & 1);
          float* o = e.bn_partial + (static_cast<long long>(split) * g.ldc + nt_mine * N_GEMM + pass * PN + et) * 3;
          o[0] = n; o[1] = mean; o[2] = fmaxf(q - s * mean, 0.f);
        }
      }
    }
  }

  tc::tc_fence_before();
  __syncthreads();
  if (CL > 1) tc::cluster_sync();  // the peer may still multicast into this CTA's stages / arrive on its barriers until it is done
  if (threadIdx.x == 0) PDL_TRACE(8);
  if (warp == 2) tc::tmem_dealloc(tmem_base, kTmemCols);
  if (threadIdx.x == 0 && *failed && e.error) *e.error = 1;
}

// ---- host side ----
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_tiled() {
  static EncodeTiledFn fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  }
  return fn;
}

// NHWC activation [N][H][W][C] viewed as a 4-D tensor (C, W, H, N); box = (32, bw, bh, bn).
int make_act_map(CUtensorMap* m, const float* base, int N, int H, int W, int C, int bw, int bh, int bn, int stride = 1) {
  EncodeTiledFn enc = encode_tiled();
  if (!enc) { pdl_set_error("cuTensorMapEncodeTiled is not available"); return -1; }
  cuuint64_t dims[4] = {(cuuint64_t)C, (cuuint64_t)W, (cuuint64_t)H, (cuuint64_t)N};
  cuuint64_t strides[3] = {(cuuint64_t)C * 4, (cuuint64_t)W * C * 4, (cuuint64_t)H * W * C * 4};
  // stride s: the box spans bw*s x bh*s input pixels and the traversal stride keeps every s-th one (bw x bh rows land in smem)
  cuuint32_t box[4] = {32, (cuuint32_t)(bw * stride), (cuuint32_t)(bh * stride), (cuuint32_t)bn};
  cuuint32_t es[4] = {1, (cuuint32_t)stride, (cuuint32_t)stride, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { pdl_set_error("cuTensorMapEncodeTiled(activation %dx%dx%dx%d, box %dx%dx%d) failed: %d", N, H, W, C, bw, bh, bn, (int)r); return -2; }
  return 0;
}

// Packed weights [rows][K] (K contiguous) as a 2-D tensor (K, rows); box = (32, box_rows).
int make_weight_map(CUtensorMap* m, const float* base, int rows, int K, int box_rows) {
  EncodeTiledFn enc = encode_tiled();
  if (!enc) { pdl_set_error("cuTensorMapEncodeTiled is not available"); return -1; }
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)K * 4};
  cuuint32_t box[2] = {32, (cuuint32_t)box_rows};
  cuuint32_t es[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                   CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { pdl_set_error("cuTensorMapEncodeTiled(weights %dx%d) failed: %d", rows, K, (int)r); return -2; }
  return 0;
}

// Dense NHWC output [N][H][W][C] for the TMA-store epilogue: box = (32, bw, bh, bn), same row order as the A box.
int make_out_map(CUtensorMap* m, float* base, int N, int H, int W, int C, int bw, int bh, int bn) {
  return make_act_map(m, base, N, H, W, C, bw, bh, bn, 1);
}
inline bool use_tma_store() {
  static int v = -1;
  if (v < 0) { const char* s = getenv("PDL_TC_TMA_STORE"); v = s ? atoi(s) : 1; }
  return v != 0;
}

// CTAs of a launch over num_tiles (= m_tiles * n_tiles) tiles: balanced waves; with fused BatchNorm statistics also a
// multiple of n_tiles, so that all tiles of a CTA share one column block (grid / n_tiles = rows of the partial buffer).
// Upper bound on the CTAs of the forward / dgrad launches that follow (0 = the whole chip); see pdl_tc_conv_max_ctas.
int g_tc_cta_cap = 0;
int tc_grid(int num_tiles, int n_tiles, bool bn_stats) {
  static const int env_cap = getenv("PDL_TC_MAX_CTAS") ? atoi(getenv("PDL_TC_MAX_CTAS")) : 0;
  const int cap = env_cap > 0 ? env_cap : g_tc_cta_cap;
  const int sms = (cap > 0 && cap < 148) ? cap : 148;
  const int waves = (num_tiles + sms - 1) / sms;
  int grid = (num_tiles + waves - 1) / waves;
  if (bn_stats) {
    grid = grid / n_tiles * n_tiles;
    if (grid < n_tiles) grid = n_tiles;
  }
  return grid;
}

inline bool cta_pairs_allowed(int num_tiles, int n_tiles, int n_gemm);
// CTAs of a launch, and whether it runs as CTA pairs (tc_conv_kernel<.., CL = 2>): the one place that decides both, shared by
// the launcher and by the functions that size the BatchNorm-statistics buffer (rows = grid / n_tiles).
inline int planned_grid(int num_tiles, int n_tiles, int n_gemm, bool bn_stats, bool* pairs_out) {
  const bool pairs = cta_pairs_allowed(num_tiles, n_tiles, n_gemm);
  if (pairs_out) *pairs_out = pairs;
  if (!pairs) return tc_grid(num_tiles, n_tiles, bn_stats);
  // pairs of CTAs step through pairs of tiles: balanced waves over 74 pairs; with fused statistics the pair count is a
  // multiple of n_tiles so that every CTA keeps one column block
  const int num_pairs = num_tiles / 2;
  const int waves = (num_pairs + 73) / 74;
  int pairs_n = (num_pairs + waves - 1) / waves;
  if (bn_stats) {
    pairs_n = pairs_n / n_tiles * n_tiles;
    if (pairs_n < n_tiles) pairs_n = n_tiles;
  }
  return 2 * pairs_n;
}

// Whether a launch runs as CTA pairs sharing the weight tile (tc_conv_kernel<.., CL = 2>): an even number of M tiles (pairs
// never straddle a column block), at least a wave of work (below that the kernels are latency-bound, and a pair must be
// co-scheduled on one TPC), and a weight tile whose halves are whole swizzle atoms.
// OPT-IN (PDL_TC_CLUSTER=1; 2 forces pairs on small launches too, which the tests use).  Measured on B200, ResNet-50 batch 64
// (profiles/tc_conv_cta_pairs.md): every layer within +-2 % of the single-CTA kernel, the step 13.15 vs 13.09 ms.  Halving the
// L2 READS of the weight tile does not help because what bounds these kernels is the bytes that LAND in each SM's shared
// memory per MAC -- a multicast tile still arrives in full in both SMs.
inline bool cta_pairs_allowed(int num_tiles, int n_tiles, int n_gemm) {
  static const int mode = getenv("PDL_TC_CLUSTER") ? atoi(getenv("PDL_TC_CLUSTER")) : 0;
  if (mode == 0 || n_gemm < 64 || n_tiles > 74) return false;
  const int m_tiles = num_tiles / n_tiles;
  return (m_tiles % 2 == 0) && (mode == 2 || num_tiles >= 148);
}

template <int N_GEMM, int EPI>
int launch_tc(const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& o, const TcGeom& g, const EpiArgs& e, cudaStream_t stream) {
  static bool configured = false;
  if (!configured) {
    PDL_CUDA_CHECK(cudaFuncSetAttribute(tc_conv_kernel<N_GEMM, EPI, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<N_GEMM>::TOTAL));
    PDL_CUDA_CHECK(cudaFuncSetAttribute(tc_conv_kernel<N_GEMM, EPI, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, Smem<N_GEMM>::TOTAL));
    configured = true;
  }
  // balanced waves: 200 tiles run as 100 CTAs x 2 tiles, not 148 CTAs of which 52 take two -- same makespan, and the
  // SMs left over go to whatever runs beside this kernel (the wgrad branch of the ConvNet graph)
  bool pairs = false;
  const int grid = planned_grid(g.num_tiles, g.n_tiles, N_GEMM, e.bn_partial != nullptr, &pairs);
  if (pairs && g.w_base != nullptr) {
    CUtensorMap bh;  // the same packed weights, box = half a column tile: each CTA of a pair fetches (and multicasts) one half
    if (int rc = make_weight_map(&bh, g.w_base, g.w_rows, g.w_K, N_GEMM / 2)) return rc;
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = dim3(grid);
    cfg.blockDim = dim3(kThreads);
    cfg.dynamicSmemBytes = Smem<N_GEMM>::TOTAL;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_dependent_launch_enabled() ? 2 : 1;
    PDL_CUDA_CHECK(cudaLaunchKernelEx(&cfg, tc_conv_kernel<N_GEMM, EPI, 2>, a, bh, o, g, e));
    PDL_LAUNCH_CHECK();
    return 0;
  }
  if (pairs) { pdl_set_error("tc_conv: a paired launch needs the weight description in its geometry"); return -3; }
  PDL_CUDA_CHECK(launch_dependent(tc_conv_kernel<N_GEMM, EPI, 1>, dim3(grid), dim3(kThreads), Smem<N_GEMM>::TOTAL, stream, a, b, o, g, e));
  PDL_LAUNCH_CHECK();
  return 0;
}

template <int EPI>
int dispatch_n(int n_gemm, const CUtensorMap& a, const CUtensorMap& b, const CUtensorMap& o, const TcGeom& g, const EpiArgs& e,
               cudaStream_t stream) {
  switch (n_gemm) {
    case 32: return launch_tc<32, EPI>(a, b, o, g, e, stream);
    case 64: return launch_tc<64, EPI>(a, b, o, g, e, stream);
    case 128: return launch_tc<128, EPI>(a, b, o, g, e, stream);
    case 256:
      if constexpr (EPI == EPI_STORE) {
        if (e.tma_store) return launch_tc<256, EPI_STORE>(a, b, o, g, e, stream);
      }
      pdl_set_error("tc_conv: 256-column tiles need the TMA-store epilogue");
      return -3;
    default: pdl_set_error("tc_conv: unsupported GEMM N %d", n_gemm); return -3;
  }
}

// Chooses the M-tile box for an output of Hout x Wout pixels per image.  Returns false if no box fits.
bool choose_box(int Hout, int Wout, int need_whole_image, TcGeom* g) {
  if (Wout > 128) return false;
  if (Hout * Wout <= 128) {
    g->BW = Wout; g->BH = Hout; g->BN = 128 / (Hout * Wout); g->tiles_per_img = 1;
    if (g->BN > 1 && !need_whole_image && false) g->BN = 1;
    return true;
  }
  if (need_whole_image) return false;
  g->BW = Wout; g->BH = 128 / Wout; g->BN = 1; g->tiles_per_img = (Hout + g->BH - 1) / g->BH;
  return g->BH >= 1;
}

__global__ void pack_weights_kernel(const float* __restrict__ w, float* __restrict__ fwd, float* __restrict__ bwd, int Cout, int Cin, int taps) {
  // w[co][ci][tap] -> fwd[tap][co][ci] (forward B operand, K = ci) and bwd[tap][ci][co] (dgrad B operand, K = co)
  int i = blockIdx.x * blockDim.x + threadIdx.x;
  int total = Cout * Cin * taps;
  if (i >= total) return;
  int tap = i % taps, ci = (i / taps) % Cin, co = i / (taps * Cin);
  float v = round_tf32(w[i]);
  if (fwd) fwd[(tap * Cout + co) * Cin + ci] = v;
  if (bwd) bwd[(tap * Cin + ci) * Cout + co] = v;
}

}  // namespace

static long long* g_trace = nullptr;
// Profiling aid: device buffer of >= 16 int64 that CTA 0 of the next conv launches fills with clock64() milestones.
PDL_API void pdl_tc_set_trace(long long* buf) { g_trace = buf; }

// 1 if the tensor-core forward path covers this shape (stride 1, Cin % 32 == 0, Cout in {32,64,128}).
// `wide`: the launch uses the TMA-store epilogue, which is the only one that handles 256-column tiles (in two passes).
static int n_gemm_for(int channels, bool wide = false, int m_tiles = 0) {  // column tile for `channels` outputs, 0 if none
  if (channels == 32 || channels == 64) return channels;
  static const int forced = getenv("PDL_TC_NGEMM") ? atoi(getenv("PDL_TC_NGEMM")) : 0;  // experiment knob (64 | 128 | 256)
  if (forced == 64 && channels % 64 == 0) return 64;
  // 256 columns cut the L2 traffic per MAC by a quarter but halve the number of tiles: measured on ResNet-50's layers
  // (batch 64) they win from ~96 tiles up (14x14 and larger maps: 0.049 -> 0.040 ms) and lose below (7x7 maps: 0.056 -> 0.069)
  if (wide && forced != 128 && channels % 256 == 0 && (forced == 256 || static_cast<long long>(m_tiles) * (channels / 256) >= 96)) return 256;
  return channels % 128 == 0 ? 128 : 0;
}
// M tiles of a launch whose box was chosen by choose_box
static int m_tiles_of(const TcGeom& g, int N) { return g.BN > 1 ? (N + g.BN - 1) / g.BN : N * g.tiles_per_img; }

// Bound the grid of the forward / dgrad launches that follow (0 lifts the bound); returns the previous value.  Same reason as
// pdl_tc_wgrad_max_ctas: a caller that wants another kernel to run BESIDE this one leaves it some SMs.
PDL_API int pdl_tc_conv_max_ctas(int cap) {
  const int old = g_tc_cta_cap;
  g_tc_cta_cap = cap < 0 ? 0 : cap;
  return old;
}

PDL_API int pdl_tc_conv_supported(int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad, int pool) {
  if ((stride != 1 && stride != 2) || (stride == 2 && pool) || Cin % 32 || !n_gemm_for(Cout) || (pool && Cout > 128)) return 0;
  int P = (H + 2 * pad - KH) / stride + 1, Q = (W + 2 * pad - KW) / stride + 1;
  if (Q * stride > 256) return 0;
  TcGeom g{};
  return choose_box(P, Q, pool, &g) ? 1 : 0;
}

namespace {
// All convolution weights of a model in ONE launch (a per-layer pack is a ~5 us latency-bound kernel, 52 of them in a
// ResNet-50 step, and its element-wise transposing writes touch one 32-byte sector per float).  A CTA owns a tile of
// 32 output x 32 input channels x taps (<= 9): the tile is read with contiguous runs of 32*taps floats per output
// channel, staged in shared memory (TF32-rounded), and written out as 128-byte rows in both operand layouts,
// fwd[tap][co][ci] and bwd[tap][ci][co].  table[t] = {w, packed_fwd, packed_bwd, Cout, Cin, taps} as int64;
// prefix[t] = first tile of tensor t.
constexpr int kPackMaxTaps = 9;
__global__ void __launch_bounds__(256) pack_weights_multi_kernel(const long long* __restrict__ table, const int* __restrict__ prefix, int num) {
  __shared__ float tile[32][32 * kPackMaxTaps + 1];  // [co][ci * taps + tap]
  int lo = 0, hi = num - 1;  // binary search: prefix[lo] <= blockIdx.x < prefix[lo + 1]
  while (lo < hi) {
    const int mid = (lo + hi + 1) / 2;
    if (prefix[mid] <= static_cast<int>(blockIdx.x)) lo = mid; else hi = mid - 1;
  }
  const long long* row = table + lo * 6;
  const float* w = reinterpret_cast<const float*>(row[0]);
  float* fwd = reinterpret_cast<float*>(row[1]);
  float* bwd = reinterpret_cast<float*>(row[2]);
  const int Cout = static_cast<int>(row[3]), Cin = static_cast<int>(row[4]), taps = static_cast<int>(row[5]);
  const int t = blockIdx.x - prefix[lo], ci_tiles = Cin / 32;
  const int co0 = (t / ci_tiles) * 32, ci0 = (t % ci_tiles) * 32;
  const int run = 32 * taps;  // contiguous floats per output channel in w: [ci0 .. ci0+32) x taps
  for (int k = threadIdx.x; k < 32 * run; k += 256) {
    const int co = k / run, e = k % run;
    tile[co][e] = round_tf32(w[(static_cast<long long>(co0 + co) * Cin + ci0) * taps + e]);
  }
  __syncthreads();
  const int lane = threadIdx.x % 32, wp = threadIdx.x / 32;
  for (int r = wp; r < 32 * taps; r += 8) {  // one 128-byte row per warp iteration, for each layout
    const int tap = r / 32, j = r % 32;
    if (fwd) fwd[(static_cast<long long>(tap) * Cout + co0 + j) * Cin + ci0 + lane] = tile[j][lane * taps + tap];   // row = (tap, co), lanes = ci
    if (bwd) bwd[(static_cast<long long>(tap) * Cin + ci0 + j) * Cout + co0 + lane] = tile[lane][j * taps + tap];   // row = (tap, ci), lanes = co
  }
}
}  // namespace

PDL_API int pdl_tc_pack_max_taps() { return kPackMaxTaps; }
// table: [num][6] int64 device rows {w, packed_fwd, packed_bwd, Cout, Cin, taps} with Cout % 32 == Cin % 32 == 0 and
// taps <= pdl_tc_pack_max_taps(); prefix: [num + 1] int32 offsets in tiles of 32 x 32 channels ((Cout/32)*(Cin/32) per tensor).
PDL_API int pdl_tc_pack_weights_multi(const void* table, const int* prefix, int num, int total_tiles, cudaStream_t stream) {
  if (num <= 0 || total_tiles <= 0) return 0;
  pack_weights_multi_kernel<<<total_tiles, 256, 0, stream>>>(static_cast<const long long*>(table), prefix, num);
  PDL_LAUNCH_CHECK();
  return 0;
}

// w[Cout][Cin][KH][KW] -> packed_fwd[tap][Cout][Cin], packed_bwd[tap][Cin][Cout] (either may be null).
PDL_API int pdl_tc_pack_weights(const float* w, float* packed_fwd, float* packed_bwd, int Cout, int Cin, int taps, cudaStream_t stream) {
  int total = Cout * Cin * taps;
  pack_weights_kernel<<<pdl_ceil_div(total, 256), 256, 0, stream>>>(w, packed_fwd, packed_bwd, Cout, Cin, taps);
  PDL_LAUNCH_CHECK();
  return 0;
}

// Forward: y = conv(x, w) + bias, optional ReLU; with pool_stride in {1,2} the 2x2 max-pool is fused and
// `y`/`idx` receive the POOLED output [N][PH][PW][Cout] and its argmax (0..3).  x is NHWC.
static int tc_fwd_impl(const float* x, const float* packed_w, const float* bias, float* y, uint8_t* idx, int N, int H, int W, int Cin, int Cout,
                       int KH, int KW, int stride, int pad, int relu, int pool_stride, int* error_flag, cudaStream_t stream,
                       float* bn_partial = nullptr) {
  TcGeom g{};
  const int P = (H + 2 * pad - KH) / stride + 1, Q = (W + 2 * pad - KW) / stride + 1;
  g.in_stride = stride;
  g.tap_r0 = g.tap_s0 = 0; g.tap_step = 1; g.S_full = KW; g.out_stride = 1; g.out_oh = g.out_ow = 0; g.Hc = P;
  if (!choose_box(P, Q, pool_stride > 0, &g)) { pdl_set_error("tc_conv fwd: no tile for %dx%d output", P, Q); return -4; }
  const int ng = n_gemm_for(Cout, pool_stride == 0 && use_tma_store(), m_tiles_of(g, N));
  if (!ng || Cin % 32 || (pool_stride > 0 && Cout > 128)) { pdl_set_error("tc_conv fwd: unsupported channels %d -> %d", Cin, Cout); return -3; }
  g.n_tiles = Cout / ng; g.ldc = Cout;
  g.num_tiles = (g.BN > 1 ? (N + g.BN - 1) / g.BN : N * g.tiles_per_img) * g.n_tiles;
  g.R = KH; g.S = KW; g.kchunks = Cin / 32; g.a_w0 = -pad; g.a_h0 = -pad; g.tap_sign = 1; g.Nimg = N; g.Hout = P; g.Wout = Q;
  CUtensorMap ta, tb;
  if (int rc = make_act_map(&ta, x, N, H, W, Cin, g.BW, g.BH, g.BN, stride)) return rc;
  if (int rc = make_weight_map(&tb, packed_w, KH * KW * Cout, Cin, ng)) return rc;
  g.w_base = packed_w; g.w_rows = KH * KW * Cout; g.w_K = Cin;
  const int tma_store = pool_stride == 0 && use_tma_store();
  CUtensorMap to = ta;
  if (tma_store) { if (int rc = make_out_map(&to, y, N, P, Q, Cout, g.BW, g.BH, g.BN)) return rc; }
  if (bn_partial && (!tma_store || bias || relu)) { pdl_set_error("tc_conv fwd: fused BatchNorm statistics need the plain TMA-store epilogue"); return -3; }
  EpiArgs e{y, idx, bias, relu, pool_stride, tma_store, error_flag, g_trace, nullptr, nullptr, bn_partial};
  return pool_stride > 0 ? dispatch_n<EPI_RELU_POOL>(ng, ta, tb, to, g, e, stream) : dispatch_n<EPI_STORE>(ng, ta, tb, to, g, e, stream);
}

// Rows of the statistics buffer pdl_tc_conv2d_fwd_bnstats fills ([splits][Cout][3] floats: count, mean, M2 per CTA group),
// or 0 when the shape is not covered by the tensor-core forward.
PDL_API int pdl_tc_conv_bn_splits(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad) {
  if (!pdl_tc_conv_supported(H, W, Cin, Cout, KH, KW, stride, pad, 0) || !use_tma_store()) return 0;
  TcGeom g{};
  const int P = (H + 2 * pad - KH) / stride + 1, Q = (W + 2 * pad - KW) / stride + 1;
  if (!choose_box(P, Q, 0, &g)) return 0;
  const int ng = n_gemm_for(Cout, true, m_tiles_of(g, N));
  const int n_tiles = Cout / ng;
  const int num_tiles = m_tiles_of(g, N) * n_tiles;
  return planned_grid(num_tiles, n_tiles, ng, true, nullptr) / n_tiles;
}

// Bias-free forward (stride 1|2) that also leaves the per-channel BatchNorm statistics of its output in bn_partial, in
// the partial format of pdl_bn_fwd_train_prestats -- the normalisation layer then skips its own pass over y.
PDL_API int pdl_tc_conv2d_fwd_bnstats(const float* x, const float* packed_w, float* y, float* bn_partial, int N, int H, int W, int Cin, int Cout,
                                      int KH, int KW, int stride, int pad, int* error_flag, cudaStream_t stream) {
  if (stride != 1 && stride != 2) { pdl_set_error("tc_conv fwd: stride %d is not supported", stride); return -3; }
  return tc_fwd_impl(x, packed_w, nullptr, y, nullptr, N, H, W, Cin, Cout, KH, KW, stride, pad, 0, 0, error_flag, stream, bn_partial);
}

PDL_API int pdl_tc_conv2d_fwd(const float* x, const float* packed_w, const float* bias, float* y, uint8_t* idx, int N, int H, int W, int Cin,
                              int Cout, int KH, int KW, int pad, int relu, int pool_stride, int* error_flag, cudaStream_t stream) {
  return tc_fwd_impl(x, packed_w, bias, y, idx, N, H, W, Cin, Cout, KH, KW, 1, pad, relu, pool_stride, error_flag, stream);
}

// Forward with a convolution stride of 1 or 2 (no fused pooling): the TMA traversal stride subsamples the input box.
PDL_API int pdl_tc_conv2d_fwd_strided(const float* x, const float* packed_w, const float* bias, float* y, int N, int H, int W, int Cin, int Cout,
                                      int KH, int KW, int stride, int pad, int relu, int* error_flag, cudaStream_t stream) {
  if (stride != 1 && stride != 2) { pdl_set_error("tc_conv fwd: stride %d is not supported", stride); return -3; }
  return tc_fwd_impl(x, packed_w, bias, y, nullptr, N, H, W, Cin, Cout, KH, KW, stride, pad, relu, 0, error_flag, stream);
}

// dgrad (stride 1): dx[N][H][W][Cin] from dy[N][P][Q][Cout] and packed_bwd[tap][Cin][Cout].
// dgrad for convolution stride 1 or 2.  Stride 2 is solved as four stride-1 problems, one per parity class
// (h % 2, w % 2) of dx: only the taps with matching parity reach a class, the class is a dense sub-grid of dx,
// and its outputs are written with a stride of 2.  Classes no tap reaches (1x1 filters) stay zero.
static int dgrad_impl(const float* dy, const float* packed_w_bwd, float* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride,
                      int pad, int* error_flag, cudaStream_t stream, const float* add_src);
PDL_API int pdl_tc_conv2d_dgrad_strided(const float* dy, const float* packed_w_bwd, float* dx, int N, int H, int W, int Cin, int Cout, int KH,
                                        int KW, int stride, int pad, int* error_flag, cudaStream_t stream) {
  return dgrad_impl(dy, packed_w_bwd, dx, N, H, W, Cin, Cout, KH, KW, stride, pad, error_flag, stream, nullptr);
}
// dx = dgrad(dy, w) + add_src (stride 1): the skip connection's gradient joins in the epilogue instead of a separate
// element-wise pass over two full tensors.  Returns -5 if the TMA-store epilogue is disabled (caller adds separately).
PDL_API int pdl_tc_conv2d_dgrad_add(const float* dy, const float* packed_w_bwd, const float* add_src, float* dx, int N, int H, int W, int Cin,
                                    int Cout, int KH, int KW, int pad, int* error_flag, cudaStream_t stream) {
  if (!use_tma_store()) { pdl_set_error("tc_conv dgrad+add needs the TMA-store epilogue"); return -5; }
  return dgrad_impl(dy, packed_w_bwd, dx, N, H, W, Cin, Cout, KH, KW, 1, pad, error_flag, stream, add_src);
}
static int dgrad_impl(const float* dy, const float* packed_w_bwd, float* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride,
                      int pad, int* error_flag, cudaStream_t stream, const float* add_src) {
  if (stride != 1 && stride != 2) { pdl_set_error("tc_conv dgrad: stride %d is not supported", stride); return -3; }
  const int P = (H + 2 * pad - KH) / stride + 1, Q = (W + 2 * pad - KW) / stride + 1;
  int m_tiles = 0;
  {
    TcGeom g0{};
    if (stride == 1 && choose_box(H, W, 0, &g0)) m_tiles = m_tiles_of(g0, N);
  }
  const int ng = n_gemm_for(Cin, stride == 1 && use_tma_store(), m_tiles);  // parity-class launches (stride 2) store directly
  if (Cout % 32 || !ng) { pdl_set_error("tc_conv dgrad: unsupported channels %d -> %d", Cout, Cin); return -3; }
  CUtensorMap tb;
  if (int rc = make_weight_map(&tb, packed_w_bwd, KH * KW * Cin, Cout, ng)) return rc;
  bool need_zero = false;
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      const int r0 = (ph + pad) % stride, s0 = (pw + pad) % stride;
      if (r0 >= KH || s0 >= KW || ph >= H || pw >= W) need_zero = true;
    }
  if (need_zero) PDL_CUDA_CHECK(cudaMemsetAsync(dx, 0, sizeof(float) * static_cast<size_t>(N) * H * W * Cin, stream));
  for (int ph = 0; ph < stride; ++ph)
    for (int pw = 0; pw < stride; ++pw) {
      // dx[stride*a + ph, stride*b + pw] = sum over taps r = r0 + stride*i, s = s0 + stride*j of dy[a + (ph+pad-r)/stride, b + (pw+pad-s)/stride] * w[r, s]
      const int r0 = (ph + pad) % stride, s0 = (pw + pad) % stride;
      if (r0 >= KH || s0 >= KW || ph >= H || pw >= W) continue;
      TcGeom g{};
      const int Hc = (H - ph + stride - 1) / stride, Wc = (W - pw + stride - 1) / stride;
      if (!choose_box(Hc, Wc, 0, &g)) { pdl_set_error("tc_conv dgrad: no tile for %dx%d", Hc, Wc); return -4; }
      g.n_tiles = Cin / ng; g.ldc = Cin;
      g.num_tiles = (g.BN > 1 ? (N + g.BN - 1) / g.BN : N * g.tiles_per_img) * g.n_tiles;
      g.in_stride = 1;
      g.R = (KH - r0 + stride - 1) / stride; g.S = (KW - s0 + stride - 1) / stride; g.kchunks = Cout / 32;
      g.tap_r0 = r0; g.tap_s0 = s0; g.tap_step = stride; g.S_full = KW;
      g.a_h0 = (ph + pad - r0) / stride; g.a_w0 = (pw + pad - s0) / stride; g.tap_sign = -1;
      g.out_stride = stride; g.out_oh = ph; g.out_ow = pw; g.Hc = Hc;
      g.Nimg = N; g.Hout = H; g.Wout = W;
      g.w_base = packed_w_bwd; g.w_rows = KH * KW * Cin; g.w_K = Cout;
      CUtensorMap ta;
      if (int rc = make_act_map(&ta, dy, N, P, Q, Cout, g.BW, g.BH, g.BN)) return rc;
      const int tma_store = stride == 1 && use_tma_store();  // parity classes scatter with a stride: direct stores
      CUtensorMap to = ta;
      if (tma_store) { if (int rc = make_out_map(&to, dx, N, H, W, Cin, g.BW, g.BH, g.BN)) return rc; }
      EpiArgs e{dx, nullptr, nullptr, 0, 0, tma_store, error_flag, g_trace, nullptr, nullptr, nullptr, tma_store ? add_src : nullptr};
      if (int rc = dispatch_n<EPI_STORE>(ng, ta, tb, to, g, e, stream)) return rc;
    }
  return 0;
}

// dgrad of a stride-1 convolution whose INPUT is a ReLU + 2x2 max-pool (stride pool_stride) of the layer below, with
// that pool/ReLU backward fused into the epilogue: instead of the gradient of the pooled tensor [N][H][W][Cin] this
// writes dz[N][ZH][ZW][Cin] (ZH = (H-1)*pool_stride + 2, TF32-rounded: it is the next dgrad/wgrad operand) and the
// per-tile column sums dbpart[pdl_tc_dgrad_poolbwd_tiles()][Cin] (bias-gradient partials).  idx / pooled are the pool's
// arg-max codes (0..3) and outputs, [N][H][W][Cin].  Needs whole images per tile (H*W <= 128).
PDL_API int pdl_tc_dgrad_poolbwd_tiles(int N, int H, int W) {
  TcGeom g{};
  if (!choose_box(H, W, 1, &g)) return 0;
  return g.BN > 1 ? (N + g.BN - 1) / g.BN : N;
}
PDL_API int pdl_tc_conv2d_dgrad_poolbwd(const float* dy, const float* packed_w_bwd, const uint8_t* idx, const float* pooled, float* dz,
                                        float* dbpart, int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad, int pool_stride,
                                        int* error_flag, cudaStream_t stream) {
  const int P = H + 2 * pad - KH + 1, Q = W + 2 * pad - KW + 1;
  const int ng = n_gemm_for(Cin);
  if (Cout % 32 || !ng || (pool_stride != 1 && pool_stride != 2)) { pdl_set_error("tc_conv dgrad+pool: unsupported channels/stride"); return -3; }
  TcGeom g{};
  if (!choose_box(H, W, 1, &g)) { pdl_set_error("tc_conv dgrad+pool: %dx%d images do not fit one tile", H, W); return -4; }
  g.n_tiles = Cin / ng; g.ldc = Cin;
  g.num_tiles = (g.BN > 1 ? (N + g.BN - 1) / g.BN : N) * g.n_tiles;
  g.in_stride = 1;
  g.R = KH; g.S = KW; g.kchunks = Cout / 32;
  g.tap_r0 = g.tap_s0 = 0; g.tap_step = 1; g.S_full = KW;
  g.a_h0 = pad; g.a_w0 = pad; g.tap_sign = -1;
  g.out_stride = 1; g.out_oh = g.out_ow = 0; g.Hc = H;
  g.Nimg = N; g.Hout = H; g.Wout = W;
  CUtensorMap ta, tb;
  if (int rc = make_weight_map(&tb, packed_w_bwd, KH * KW * Cin, Cout, ng)) return rc;
  g.w_base = packed_w_bwd; g.w_rows = KH * KW * Cin; g.w_K = Cout;
  if (int rc = make_act_map(&ta, dy, N, P, Q, Cout, g.BW, g.BH, g.BN)) return rc;
  EpiArgs e{dz, const_cast<uint8_t*>(idx), nullptr, 0, pool_stride, 0, error_flag, g_trace, pooled, dbpart};
  return dispatch_n<EPI_POOL_BWD>(ng, ta, tb, ta, g, e, stream);
}

// dgrad (stride 1): dx[N][H][W][Cin] from dy[N][P][Q][Cout] and packed_bwd[tap][Cin][Cout].
PDL_API int pdl_tc_conv2d_dgrad(const float* dy, const float* packed_w_bwd, float* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                                int pad, int* error_flag, cudaStream_t stream) {
  return pdl_tc_conv2d_dgrad_strided(dy, packed_w_bwd, dx, N, H, W, Cin, Cout, KH, KW, 1, pad, error_flag, stream);
}

// ---------------------------------------------------------------------------------------------------------
// Small-Cin ("stem") convolutions on the tensor cores: the 7x7/2 3-channel first layer of the ResNets
// (reference: torchvision resnet18's conv1, example_launch.py:26).  See tc::stem_dims for the layout trick.
// ---------------------------------------------------------------------------------------------------------
namespace {

// x (any strides, element units) -> xp[N][Hp][Wp][4], zero halo and zero 4th channel, TF32-rounded
__global__ void __launch_bounds__(256) stem_pad_kernel(const float* __restrict__ x, long long sn, long long sc, long long sh, long long sw,
                                                       float4* __restrict__ xp, unsigned total, int C, int H, int W, int Hp, int Wp, int pad) {
  const unsigned i = blockIdx.x * 256u + threadIdx.x;
  if (i >= total) return;
  const int w = i % Wp;
  unsigned t = i / Wp;
  const int h = t % Hp;
  const unsigned n = t / Hp;
  const int ih = h - pad, iw = w - pad;
  float v[4] = {0.f, 0.f, 0.f, 0.f};
  if (ih >= 0 && ih < H && iw >= 0 && iw < W) {
    const float* src = x + n * sn + ih * sh + iw * sw;
#pragma unroll
    for (int c = 0; c < 4; ++c)
      if (c < C) v[c] = round_tf32(__ldg(src + c * sc));
  }
  xp[i] = make_float4(v[0], v[1], v[2], v[3]);
}

// w[co][c][r][s] -> pk[r][co][s*4 + c] (zero where s >= KW or c >= Cin), TF32-rounded
__global__ void stem_pack_weights_kernel(const float* __restrict__ w, float* __restrict__ pk, int Cout, int Cin, int KH, int KW) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= KH * Cout * 32) return;
  const int j = i % 32, co = (i / 32) % Cout, r = i / (32 * Cout);
  const int s = j / 4, c = j % 4;
  pk[i] = (s < KW && c < Cin) ? round_tf32(w[((co * Cin + c) * KH + r) * KW + s]) : 0.f;
}

}  // namespace

PDL_API int pdl_tc_stem_supported(int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad) {
  if (Cin < 1 || Cin > 4 || KW > 8 || KH > 8 || stride < 1 || stride > 2 || !n_gemm_for(Cout)) return 0;
  const tc::StemDims d = tc::stem_dims(H, W, KH, KW, stride, pad);
  TcGeom g{};
  return d.P >= 1 && d.Q >= 1 && choose_box(d.P, d.Q, 0, &g) ? 1 : 0;
}

// floats of the padded NHWC4 copy of the input
PDL_API long long pdl_tc_stem_pad_floats(int N, int H, int W, int KH, int KW, int stride, int pad) {
  const tc::StemDims d = tc::stem_dims(H, W, KH, KW, stride, pad);
  return 4ll * N * d.Hp * d.Wp;
}

// x with element strides (sn, sc, sh, sw) -> xp (pdl_tc_stem_pad_floats floats)
PDL_API int pdl_tc_stem_prepare(const float* x, long long sn, long long sc, long long sh, long long sw, float* xp, int N, int Cin, int H, int W,
                                int KH, int KW, int stride, int pad, cudaStream_t stream) {
  const tc::StemDims d = tc::stem_dims(H, W, KH, KW, stride, pad);
  const long long total = static_cast<long long>(N) * d.Hp * d.Wp;
  if (total >= (1ll << 31) || Cin > 4) { pdl_set_error("stem conv: input too large or more than 4 channels"); return -3; }
  stem_pad_kernel<<<pdl_ceil_div(total, 256), 256, 0, stream>>>(x, sn, sc, sh, sw, reinterpret_cast<float4*>(xp), static_cast<unsigned>(total), Cin,
                                                               H, W, d.Hp, d.Wp, pad);
  PDL_LAUNCH_CHECK();
  return 0;
}

// w[Cout][Cin][KH][KW] -> packed[KH][Cout][32]
PDL_API int pdl_tc_stem_pack_weights(const float* w, float* packed, int Cout, int Cin, int KH, int KW, cudaStream_t stream) {
  stem_pack_weights_kernel<<<pdl_ceil_div(KH * Cout * 32, 256), 256, 0, stream>>>(w, packed, Cout, Cin, KH, KW);
  PDL_LAUNCH_CHECK();
  return 0;
}

// y[N][P][Q][Cout] = conv(x, w) (+ bias, ReLU) from the padded copy and the packed weights.
static int stem_fwd_impl(const float* xp, const float* packed_w, const float* bias, float* y, int N, int H, int W, int Cout, int KH, int KW,
                         int stride, int pad, int relu, int* error_flag, cudaStream_t stream, float* bn_partial) {
  const tc::StemDims d = tc::stem_dims(H, W, KH, KW, stride, pad);
  TcGeom g{};
  if (!choose_box(d.P, d.Q, 0, &g)) { pdl_set_error("stem conv: no tile for %dx%d output", d.P, d.Q); return -4; }
  const int ng = n_gemm_for(Cout);
  if (!ng) { pdl_set_error("stem conv: unsupported output channels %d", Cout); return -3; }
  g.in_stride = stride;
  g.tap_r0 = g.tap_s0 = 0; g.tap_step = 1; g.S_full = 1; g.out_stride = 1; g.out_oh = g.out_ow = 0; g.Hc = d.P;
  g.n_tiles = Cout / ng; g.ldc = Cout;
  g.num_tiles = (g.BN > 1 ? (N + g.BN - 1) / g.BN : N * g.tiles_per_img) * g.n_tiles;
  g.R = KH; g.S = 1; g.kchunks = 1; g.a_w0 = 0; g.a_h0 = 0; g.tap_sign = 1; g.Nimg = N; g.Hout = d.P; g.Wout = d.Q;
  EncodeTiledFn enc = encode_tiled();
  if (!enc) { pdl_set_error("cuTensorMapEncodeTiled is not available"); return -1; }
  CUtensorMap ta, tb;
  {
    // virtual view (32 floats, Q windows, Hp rows, N): window q starts `stride` pixels after window q-1
    cuuint64_t dims[4] = {32, (cuuint64_t)d.Q, (cuuint64_t)d.Hp, (cuuint64_t)N};
    cuuint64_t strides[3] = {(cuuint64_t)16 * stride, (cuuint64_t)d.Wp * 16, (cuuint64_t)d.Hp * d.Wp * 16};
    cuuint32_t box[4] = {32, (cuuint32_t)g.BW, (cuuint32_t)(g.BH * stride), (cuuint32_t)g.BN};
    cuuint32_t es[4] = {1, 1, (cuuint32_t)stride, 1};
    CUresult r = enc(&ta, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, const_cast<float*>(xp), dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { pdl_set_error("cuTensorMapEncodeTiled(stem input, overlapping windows) failed: %d", (int)r); return -2; }
  }
  if (int rc = make_weight_map(&tb, packed_w, KH * Cout, 32, ng)) return rc;
  g.w_base = packed_w; g.w_rows = KH * Cout; g.w_K = 32;
  const int tma_store = use_tma_store();
  CUtensorMap to = ta;
  if (tma_store) { if (int rc = make_out_map(&to, y, N, d.P, d.Q, Cout, g.BW, g.BH, g.BN)) return rc; }
  if (bn_partial && (!tma_store || bias || relu)) { pdl_set_error("stem conv: fused BatchNorm statistics need the plain TMA-store epilogue"); return -3; }
  EpiArgs e{y, nullptr, bias, relu, 0, tma_store, error_flag, g_trace, nullptr, nullptr, bn_partial};
  return dispatch_n<EPI_STORE>(ng, ta, tb, to, g, e, stream);
}
PDL_API int pdl_tc_stem_fwd(const float* xp, const float* packed_w, const float* bias, float* y, int N, int H, int W, int Cout, int KH, int KW,
                            int stride, int pad, int relu, int* error_flag, cudaStream_t stream) {
  return stem_fwd_impl(xp, packed_w, bias, y, N, H, W, Cout, KH, KW, stride, pad, relu, error_flag, stream, nullptr);
}
// Stem forward that also produces the BatchNorm statistics of its output (see pdl_tc_conv2d_fwd_bnstats).
PDL_API int pdl_tc_stem_bn_splits(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad) {
  if (!pdl_tc_stem_supported(H, W, Cin, Cout, KH, KW, stride, pad) || !use_tma_store()) return 0;
  const tc::StemDims d = tc::stem_dims(H, W, KH, KW, stride, pad);
  TcGeom g{};
  if (!choose_box(d.P, d.Q, 0, &g)) return 0;
  const int ng = n_gemm_for(Cout);
  const int n_tiles = Cout / ng;
  const int num_tiles = (g.BN > 1 ? (N + g.BN - 1) / g.BN : N * g.tiles_per_img) * n_tiles;
  return planned_grid(num_tiles, n_tiles, ng, true, nullptr) / n_tiles;
}
PDL_API int pdl_tc_stem_fwd_bnstats(const float* xp, const float* packed_w, float* y, float* bn_partial, int N, int H, int W, int Cout, int KH,
                                    int KW, int stride, int pad, int* error_flag, cudaStream_t stream) {
  return stem_fwd_impl(xp, packed_w, nullptr, y, N, H, W, Cout, KH, KW, stride, pad, 0, error_flag, stream, bn_partial);
}
