// Gradient-bucket allreduce over NVLink peer memory, fused with 1/W averaging and the SGD update.
//
// Replaces the reference's hot path: DDP bucket -> ncclAllReduce -> copy-out -> optimizer.step()
// (launch_dist.py:91-92).  Every rank runs the SAME kernel; kernels talk to each
// other only through symmetric memory (csrc/cuda/symm.cu):
//   in[q]   rank q's raw gradient bucket            (peer-mapped, read over NVLink)
//   out[q]  rank q's averaged-gradient bucket       (peer-mapped, written over NVLink)
//   pad[q]  rank q's signal pad                     (monotone epoch flags, release/acquire .sys)
// Variants (all fp32 SUM, all give bit-identical results on every rank):
//   one-shot  every rank reads all W buckets and sums in rank order 0..W-1     1 hop, (W-1)*S in
//   two-shot  rank r reduces slice r, pushes the averaged slice to all peers   2 hops, 2(W-1)/W*S
//   ring      2(W-1) neighbour steps through a symmetric staging buffer        the textbook ring
//   nvls      multimem.ld_reduce on own slice + multimem.st broadcast          in-switch reduction
//   push      every rank STORES its bucket into each peer's inbox, one flag round,     lowest latency: no remote load
//             then sums the W rows of its LOCAL inbox in rank order                    is ever waited for; (W-1)*S out
// Block b of every rank handles the same element range and only synchronises with block b of
// its peers, so no grid-wide barrier is needed.  After the averaged gradient is final the same
// threads apply SGD (weight decay, momentum, nesterov) to the local parameters.
#include "comm.cuh"

namespace {

using namespace pdl_comm;
constexpr int kThreads = 512;

struct ARParams {
  const float* in[kMaxWorld];   // every rank's raw bucket (own entry = local pointer)
  float* out[kMaxWorld];        // every rank's averaged bucket
  float* ring[kMaxWorld];       // every rank's ring staging buffer (ring only)
  uint32_t* pad[kMaxWorld];     // every rank's signal pad
  const float* mc_in;           // multicast alias of `in`  (nvls only)
  float* mc_out;                // multicast alias of `out` (nvls only)
  float* param;                 // local parameters, same layout as the bucket (nullptr = no SGD)
  float* mom;                   // local momentum buffer (nullptr = momentum 0)
  SgdHyper* hyper;              // device-resident hyper-parameters
  CommState* state;
  long long n4;                 // bucket length in float4
  long long row4;               // push only: distance between two inbox rows in float4 (>= n4; fixed per inbox)
  int rank, world, channel;
  float scale;                  // 1/W for gradient averaging, 1 for a plain SUM
  unsigned long long timeout_ns;
};

__device__ __forceinline__ float4 mc_ld_reduce_f4(const float* base, long long i) {
  float4 v;
  asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
               : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(base + 4 * i) : "memory");
  return v;
}
__device__ __forceinline__ void mc_st_f4(float* base, long long i, float4 v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(base + 4 * i), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ float4 add4(float4 a, float4 b) { return make_float4(a.x + b.x, a.y + b.y, a.z + b.z, a.w + b.w); }
__device__ __forceinline__ float4 scale4(float4 a, float s) { return make_float4(a.x * s, a.y * s, a.z * s, a.w * s); }

// Tell `dst` ranks "block `blk` of rank `p.rank` reached value v".  bar.sync orders every thread's earlier
// memory operations before the signalling thread's st.release.sys, and release is cumulative in the PTX
// memory model, so the CTA's prior (local and remote) writes are visible system-wide to whoever acquires
// the flag.  (No per-thread fence.sc.sys: 512 of those per barrier cost microseconds on NVLink.)
__device__ __forceinline__ void signal_peers(const ARParams& p, int blk, uint32_t v, int only = -1) {
  __syncthreads();
  int t = threadIdx.x;
  if (t < p.world && t != p.rank && (only < 0 || t == only)) st_release_sys(p.pad[t] + flag_index(p.channel, blk, p.rank), v);
}
// Wait until the listed source ranks signalled value >= v.  Returns false (to every thread of the CTA) when a peer
// did not show up within the timeout: the caller must then neither publish results nor touch the parameters --
// whatever it would read from that peer is stale.
__device__ __forceinline__ bool wait_peers(const ARParams& p, int blk, uint32_t v, int only = -1) {
  int t = threadIdx.x;
  int failed = 0;
  if (t < p.world && t != p.rank && (only < 0 || t == only))
    failed = poll_flag(p.pad[p.rank] + flag_index(p.channel, blk, t), v, p.timeout_ns, p.state) ? 0 : 1;
  return __syncthreads_or(failed) == 0;
}
#define WAIT_OR_BAIL(...)                                   \
  if (!wait_peers(__VA_ARGS__)) {                           \
    retire(p, nblocks, 0, false);                           \
    return;                                                 \
  }

template <bool kSgd>
__device__ __forceinline__ void finish_element(const ARParams& p, long long i, float4 avg, const SgdHyper& h, bool has_mom) {
  if (kSgd) {
    float4 w = reinterpret_cast<const float4*>(p.param)[i];
    float4 m = has_mom ? reinterpret_cast<const float4*>(p.mom)[i] : make_float4(0, 0, 0, 0);
    w.x = sgd_update(w.x, avg.x, &m.x, h, has_mom);
    w.y = sgd_update(w.y, avg.y, &m.y, h, has_mom);
    w.z = sgd_update(w.z, avg.z, &m.z, h, has_mom);
    w.w = sgd_update(w.w, avg.w, &m.w, h, has_mom);
    reinterpret_cast<float4*>(p.param)[i] = w;
    if (has_mom) reinterpret_cast<float4*>(p.mom)[i] = m;
  }
}

// The last block of the launch advances the channel epoch (all blocks have read it by then).
__device__ __forceinline__ void retire(const ARParams& p, int nblocks, uint32_t epoch_advance, bool sgd) {
  __syncthreads();
  if (threadIdx.x == 0) {
    __threadfence();
    uint32_t done = atomicAdd(&p.state->finished[p.channel], 1u);
    if (done == static_cast<uint32_t>(nblocks) - 1) {
      p.state->finished[p.channel] = 0;
      p.state->epoch[p.channel] += epoch_advance;
      if (sgd) p.hyper->first_step = 0;
      __threadfence();
    }
  }
}

enum Algo : int { ONESHOT = 0, TWOSHOT = 1, RING = 2, NVLS = 3, PUSH = 4 };

// NVLink round trips are ~2-3 us: a thread must have several independent peer loads in flight.  The loads are
// volatile asm (program order is kept), so every load of a batch is issued BEFORE the first dependent store.
template <int U, typename LoadF, typename UseF>
__device__ __forceinline__ void batched(long long lo, long long hi, LoadF load, UseF use) {
  for (long long base = lo + threadIdx.x; base < hi; base += static_cast<long long>(kThreads) * U) {
    float4 v[U];
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = base + static_cast<long long>(u) * kThreads;
      if (i < hi) v[u] = load(i);
    }
#pragma unroll
    for (int u = 0; u < U; ++u) {
      const long long i = base + static_cast<long long>(u) * kThreads;
      if (i < hi) use(i, v[u]);
    }
  }
}

template <int kAlgo, bool kSgd>
__device__ void allreduce_body(const ARParams& p, int blk, int nblocks) {
  const int W = p.world, r = p.rank, tid = threadIdx.x;
  const uint32_t e = p.state->epoch[p.channel];
  const float inv = p.scale;
  if (__syncthreads_or(p.state->error != 0)) {  // a previous collective of this communicator lost a peer: leave gradients and parameters untouched
    retire(p, nblocks, 0, false);
    return;
  }
  SgdHyper h{};
  bool has_mom = false;
  if (kSgd) { h = *p.hyper; has_mom = p.mom != nullptr && h.momentum != 0.0f; }

  if (W == 1) {
    long long per = (p.n4 + nblocks - 1) / nblocks, lo = per * blk, hi = min(p.n4, lo + per);
    batched<4>(lo, hi, [&](long long i) { return scale4(reinterpret_cast<const float4*>(p.in[0])[i], inv); },
               [&](long long i, float4 v) {
                 reinterpret_cast<float4*>(p.out[0])[i] = v;
                 finish_element<kSgd>(p, i, v, h, has_mom);
               });
    retire(p, nblocks, 0, kSgd);
    return;
  }

  auto sum_peers = [&](long long i) {  // fixed rank order 0..W-1: every rank computes the same bits
    float4 acc = ld_f4(p.in[0], i);
#pragma unroll
    for (int q = 1; q < kMaxWorld; ++q)
      if (q < W) acc = add4(acc, ld_f4(p.in[q], i));
    return scale4(acc, inv);
  };
  auto sgd_use = [&](long long i, float4 v) { finish_element<kSgd>(p, i, v, h, has_mom); };

  if (kAlgo == ONESHOT) {
    long long per = (p.n4 + nblocks - 1) / nblocks, lo = per * blk, hi = min(p.n4, lo + per);
    signal_peers(p, blk, e + 1);
    WAIT_OR_BAIL(p, blk, e + 1)
    auto use = [&](long long i, float4 acc) {
      reinterpret_cast<float4*>(p.out[r])[i] = acc;
      finish_element<kSgd>(p, i, acc, h, has_mom);
    };
    if (W <= 4) batched<4>(lo, hi, sum_peers, use); else batched<2>(lo, hi, sum_peers, use);
    signal_peers(p, blk, e + 2);  // "I no longer read your bucket"
    WAIT_OR_BAIL(p, blk, e + 2)
    retire(p, nblocks, 2, kSgd);
    return;
  }

  if (kAlgo == PUSH) {
    // p.ring[q] = rank q's inbox for THIS bucket: [2 parities][W source ranks][n4] float4.  One flag round per launch is
    // enough because rows alternate between two parities: a rank overwrites parity e&1 again only in launch e+2, after it has
    // seen every peer's flag of launch e+1 -- raised by a kernel that is stream-ordered after that peer's reads of launch e.
    long long per = (p.n4 + nblocks - 1) / nblocks, lo = per * blk, hi = min(p.n4, lo + per);
    const long long row = (static_cast<long long>(e & 1u) * W + r) * p.row4;
    const float4* mine = reinterpret_cast<const float4*>(p.in[r]);
    batched<4>(lo, hi, [&](long long i) { return mine[i]; },
               [&](long long i, float4 v) {
#pragma unroll
                 for (int q = 0; q < kMaxWorld; ++q)
                   if (q < W && q != r) st_f4(p.ring[q], row + i, v);
               });
    signal_peers(p, blk, e + 1);
    WAIT_OR_BAIL(p, blk, e + 1)
    const float* inbox = p.ring[r] + 4 * (static_cast<long long>(e & 1u) * W * p.row4);
    auto sum_rows = [&](long long i) {  // fixed rank order 0..W-1: every rank computes the same bits
      float4 acc = r == 0 ? mine[i] : ld_f4(inbox, i);
#pragma unroll
      for (int q = 1; q < kMaxWorld; ++q)
        if (q < W) acc = add4(acc, q == r ? mine[i] : ld_f4(inbox, q * p.row4 + i));
      return scale4(acc, inv);
    };
    batched<2>(lo, hi, sum_rows, [&](long long i, float4 acc) {
      reinterpret_cast<float4*>(p.out[r])[i] = acc;
      finish_element<kSgd>(p, i, acc, h, has_mom);
    });
    retire(p, nblocks, 1, kSgd);
    return;
  }

  // slice layout shared by two-shot / nvls / ring: W slices of S float4; block b owns sub-range b of each slice
  const long long S = (p.n4 + W - 1) / W;
  const long long C = (S + nblocks - 1) / nblocks;
  auto sub_lo = [&](int s) { return min(p.n4, s * S + min(S, C * blk)); };
  auto sub_hi = [&](int s) { return min(p.n4, s * S + min(S, C * (blk + 1))); };

  if (kAlgo == TWOSHOT || kAlgo == NVLS) {
    signal_peers(p, blk, e + 1);
    WAIT_OR_BAIL(p, blk, e + 1)
    if (kAlgo == NVLS) {
      batched<4>(sub_lo(r), sub_hi(r), [&](long long i) { return scale4(mc_ld_reduce_f4(p.mc_in, i), inv); },
                 [&](long long i, float4 acc) { mc_st_f4(p.mc_out, i, acc); });
    } else {
      auto push = [&](long long i, float4 acc) {
#pragma unroll
        for (int q = 0; q < kMaxWorld; ++q)
          if (q < W) st_f4(p.out[q], i, acc);
      };
      if (W <= 4) batched<4>(sub_lo(r), sub_hi(r), sum_peers, push); else batched<2>(sub_lo(r), sub_hi(r), sum_peers, push);
    }
    signal_peers(p, blk, e + 2);  // my slice is in everyone's `out`; I am done reading `in`
    WAIT_OR_BAIL(p, blk, e + 2)
    if (kSgd) {
      for (int s = 0; s < W; ++s) batched<4>(sub_lo(s), sub_hi(s), [&](long long i) { return ld_f4(p.out[r], i); }, sgd_use);
    }
    retire(p, nblocks, 2, kSgd);
    return;
  }

  if (kAlgo == RING) {
    const int next = (r + 1) % W, prev = (r + W - 1) % W;
    // reduce-scatter: step s sends the running sum of slice (r-s) to the next rank's staging buffer
    for (int s = 0; s < W - 1; ++s) {
      int idx = ((r - s) % W + W) % W;
      batched<4>(sub_lo(idx), sub_hi(idx),
                 [&](long long i) {
                   float4 v = ld_f4(p.in[r], i);
                   if (s > 0) v = add4(ld_f4(p.ring[r], i), v);
                   return v;
                 },
                 [&](long long i, float4 v) { st_f4(p.ring[next], i, v); });
      signal_peers(p, blk, e + 1 + s, next);
      WAIT_OR_BAIL(p, blk, e + 1 + s, prev)
    }
    // slice (r+1) is now complete here: average it, keep it, start the all-gather
    {
      int idx = (r + 1) % W;
      batched<4>(sub_lo(idx), sub_hi(idx), [&](long long i) { return scale4(add4(ld_f4(p.ring[r], i), ld_f4(p.in[r], i)), inv); },
                 [&](long long i, float4 v) {
                   st_f4(p.out[r], i, v);
                   st_f4(p.out[next], i, v);
                 });
      signal_peers(p, blk, e + W, next);
    }
    for (int t = 1; t < W - 1; ++t) {
      WAIT_OR_BAIL(p, blk, e + W + t - 1, prev)
      int idx = ((r + 1 - t) % W + W) % W;
      batched<4>(sub_lo(idx), sub_hi(idx), [&](long long i) { return ld_f4(p.out[r], i); }, [&](long long i, float4 v) { st_f4(p.out[next], i, v); });
      signal_peers(p, blk, e + W + t, next);
    }
    WAIT_OR_BAIL(p, blk, e + 2 * W - 2, prev)
    signal_peers(p, blk, e + 2 * W - 1);  // full barrier: staging buffers may be reused
    WAIT_OR_BAIL(p, blk, e + 2 * W - 1)
    if (kSgd) {
      for (int s = 0; s < W; ++s) batched<4>(sub_lo(s), sub_hi(s), [&](long long i) { return ld_f4(p.out[r], i); }, sgd_use);
    }
    retire(p, nblocks, 2 * W - 1, kSgd);
    return;
  }
}

template <int kAlgo, bool kSgd>
__global__ void __launch_bounds__(kThreads) allreduce_kernel(ARParams p) {
  allreduce_body<kAlgo, kSgd>(p, blockIdx.x, gridDim.x);
}

// Single-GPU emulation of W ranks for tests: one cooperative grid of W*G blocks, "rank" = blockIdx / G.
struct EmuParams { ARParams r[kMaxWorld]; int blocks_per_rank; };
template <int kAlgo, bool kSgd>
__global__ void __launch_bounds__(kThreads) allreduce_emulated_kernel(EmuParams ep) {
  int vr = blockIdx.x / ep.blocks_per_rank, blk = blockIdx.x % ep.blocks_per_rank;
  allreduce_body<kAlgo, kSgd>(ep.r[vr], blk, ep.blocks_per_rank);
}

template <typename K, typename P>
int launch_algo(K kernel, const P& p, int blocks, cudaStream_t stream, bool cooperative) {
  if (cooperative) {
    void* args[] = {const_cast<P*>(&p)};
    PDL_CUDA_CHECK(cudaLaunchCooperativeKernel(reinterpret_cast<void*>(kernel), dim3(blocks), dim3(kThreads), args, 0, stream));
  } else {
    kernel<<<blocks, kThreads, 0, stream>>>(p);
    PDL_LAUNCH_CHECK();
  }
  return 0;
}

#define DISPATCH_ALGO(KERNEL, P, BLOCKS, COOP)                                                         \
  switch (algo * 2 + (sgd ? 1 : 0)) {                                                                  \
    case 0: return launch_algo(KERNEL<ONESHOT, false>, P, BLOCKS, stream, COOP);                       \
    case 1: return launch_algo(KERNEL<ONESHOT, true>, P, BLOCKS, stream, COOP);                        \
    case 2: return launch_algo(KERNEL<TWOSHOT, false>, P, BLOCKS, stream, COOP);                       \
    case 3: return launch_algo(KERNEL<TWOSHOT, true>, P, BLOCKS, stream, COOP);                        \
    case 4: return launch_algo(KERNEL<RING, false>, P, BLOCKS, stream, COOP);                          \
    case 5: return launch_algo(KERNEL<RING, true>, P, BLOCKS, stream, COOP);                           \
    case 6: return launch_algo(KERNEL<NVLS, false>, P, BLOCKS, stream, COOP);                          \
    case 7: return launch_algo(KERNEL<NVLS, true>, P, BLOCKS, stream, COOP);                           \
    case 8: return launch_algo(KERNEL<PUSH, false>, P, BLOCKS, stream, COOP);                          \
    case 9: return launch_algo(KERNEL<PUSH, true>, P, BLOCKS, stream, COOP);                           \
    default: pdl_set_error("unknown allreduce algorithm %d", algo); return -1;                         \
  }

// ---- generic peer-memory collectives for the process group (broadcast / allgather / barrier) ----
struct CopyParams {
  const float4* src[kMaxWorld];  // peers' staging buffers
  float4* dst;                   // local destination
  uint32_t* pad[kMaxWorld];
  CommState* state;
  long long n16;                 // 16-byte units per source
  int rank, world, channel, root;  // root < 0: allgather (dst holds world blocks)
  unsigned long long timeout_ns;
};

__global__ void __launch_bounds__(kThreads) pull_kernel(CopyParams c) {
  ARParams p{};  // reuse the signalling helpers
  for (int q = 0; q < kMaxWorld; ++q) p.pad[q] = c.pad[q];
  p.state = c.state; p.rank = c.rank; p.world = c.world; p.channel = c.channel; p.timeout_ns = c.timeout_ns;
  const int blk = blockIdx.x, nblocks = gridDim.x;
  const uint32_t e = c.state->epoch[c.channel];
  bool alive = __syncthreads_or(c.state->error != 0) == 0;
  if (c.world > 1 && alive) { signal_peers(p, blk, e + 1); alive = wait_peers(p, blk, e + 1); }
  if (!alive) { retire(p, nblocks, 0, false); return; }
  long long per = (c.n16 + nblocks - 1) / nblocks, lo = per * blk, hi = min(c.n16, lo + per);
  if (c.root >= 0) {
    if (c.rank != c.root)
      for (long long i = lo + threadIdx.x; i < hi; i += kThreads) c.dst[i] = c.src[c.root][i];
  } else {
    for (int q = 0; q < c.world; ++q)
      for (long long i = lo + threadIdx.x; i < hi; i += kThreads) c.dst[q * c.n16 + i] = c.src[q][i];
  }
  if (c.world > 1) { signal_peers(p, blk, e + 2); alive = wait_peers(p, blk, e + 2); }
  retire(p, nblocks, (c.world > 1 && alive) ? 2 : 0, false);
}

}  // namespace

// ---- C ABI -------------------------------------------------------------------------------------
PDL_API int pdl_comm_state_bytes() { return static_cast<int>(sizeof(CommState)); }
PDL_API int pdl_comm_state_error_offset() { return static_cast<int>(offsetof(CommState, error)); }
PDL_API int pdl_comm_state_host_error_offset() { return static_cast<int>(offsetof(CommState, host_error)); }
PDL_API int pdl_signal_pad_bytes() { return static_cast<int>(flag_index(kMaxChannels, 0, 0) * sizeof(uint32_t)); }
PDL_API int pdl_comm_max_blocks() { return kMaxBlocks; }
PDL_API int pdl_comm_max_channels() { return kMaxChannels; }
PDL_API int pdl_sgd_hyper_bytes() { return static_cast<int>(sizeof(SgdHyper)); }

// Number of CTAs for a bucket of n4 float4: enough loads in flight to cover NVLink latency, few
// enough that the flag traffic stays negligible.  Must be a pure function of (n4, world).
PDL_API int pdl_allreduce_blocks(long long n4, int world, int algo) {
  long long per_block = 1024;  // float4 per CTA: two per thread, all of a thread's peer loads in flight at once
  long long b = (n4 + per_block - 1) / per_block;
  if (b < 1) b = 1;
  int cap = 64;  // the default rule; the measured table (parallel/allreduce_table.json) may ask for up to kMaxBlocks
  if (b > cap) b = cap;
  (void)world;
  return static_cast<int>(b);
}

// Launch one bucket allreduce (+ optional fused SGD) on `stream`.  (pdl_allreduce_launch_ex: the same with an explicit inbox
// row distance for the push variant, for inboxes shared by launches of different lengths.)
//  in/out/ring/pad: arrays of `world` peer-mapped base pointers (already offset to the bucket / pad).
//  ring: the ring variant's staging buffer (n4 float4 per rank); for the push variant the bucket's inbox,
//  pdl_allreduce_push_inbox_bytes(n4, world) per rank, used by no other bucket.
PDL_API long long pdl_allreduce_push_inbox_bytes(long long n4, int world) { return 2ll * world * n4 * 16; }

PDL_API int pdl_allreduce_launch_ex(int algo, int rank, int world, int channel, long long n4, long long inbox_row4, const float* const* in,
                                    float* const* out, float* const* ring, uint32_t* const* pad, const float* mc_in, float* mc_out, float* param,
                                    float* mom, void* hyper, void* state, int blocks, double scale, double timeout_s, cudaStream_t stream) {
  if (world > kMaxWorld) { pdl_set_error("allreduce: world %d > %d", world, kMaxWorld); return -1; }
  if (channel < 0 || channel >= kMaxChannels) { pdl_set_error("allreduce: bad channel %d", channel); return -1; }
  if (blocks < 1 || blocks > kMaxBlocks) { pdl_set_error("allreduce: bad block count %d", blocks); return -1; }
  if (algo == NVLS && (!mc_in || !mc_out)) { pdl_set_error("allreduce: nvls needs multicast pointers"); return -1; }
  if ((algo == RING || algo == PUSH) && world > 1 && !ring) { pdl_set_error("allreduce: ring/push need their staging buffers"); return -1; }
  ARParams p{};
  for (int q = 0; q < world; ++q) {
    p.in[q] = in[q]; p.out[q] = out[q]; p.pad[q] = pad ? pad[q] : nullptr; p.ring[q] = ring ? ring[q] : nullptr;
  }
  p.mc_in = mc_in; p.mc_out = mc_out; p.param = param; p.mom = mom;
  p.hyper = static_cast<SgdHyper*>(hyper); p.state = static_cast<CommState*>(state);
  if (algo == PUSH && inbox_row4 < n4) { pdl_set_error("allreduce: push inbox rows (%lld float4) shorter than the bucket (%lld)", inbox_row4, n4); return -1; }
  p.n4 = n4; p.row4 = inbox_row4; p.rank = rank; p.world = world; p.channel = channel; p.scale = static_cast<float>(scale);
  p.timeout_ns = static_cast<unsigned long long>(timeout_s * 1e9);
  const bool sgd = param != nullptr;
  DISPATCH_ALGO(allreduce_kernel, p, blocks, false)
}

PDL_API int pdl_allreduce_launch(int algo, int rank, int world, int channel, long long n4, const float* const* in, float* const* out,
                                 float* const* ring, uint32_t* const* pad, const float* mc_in, float* mc_out, float* param, float* mom,
                                 void* hyper, void* state, int blocks, double scale, double timeout_s, cudaStream_t stream) {
  return pdl_allreduce_launch_ex(algo, rank, world, channel, n4, n4, in, out, ring, pad, mc_in, mc_out, param, mom, hyper, state, blocks, scale,
                                 timeout_s, stream);
}

// ---- one call per bucket: the reducer's launch path (reference: Reducer::mark_bucket_ready -> all_reduce_bucket on the
// process group's stream) -------------------------------------------------------------------------------
// A handle keeps everything that is fixed for a bucket (peer pointers, variant, CTA count, two events).  pdl_bucket_reduce
// then does what the autograd hook needs when the bucket's last gradient has arrived, in ONE native call:
//   record `ready` on the compute stream -> make the communication stream wait for it -> launch the fused kernel there ->
//   record `done` on the communication stream.   pdl_bucket_wait makes a stream wait for `done` (end of backward).
// Works under stream capture (the event edges become graph dependencies).
namespace {
struct BucketHandle {
  ARParams p;
  int algo, blocks;
  cudaEvent_t ready, done;
};
}  // namespace

PDL_API int pdl_bucket_create(int algo, int rank, int world, int channel, long long n4, long long inbox_row4, const float* const* in,
                              float* const* out, float* const* ring, uint32_t* const* pad, const float* mc_in, float* mc_out, void* state,
                              int blocks, double scale, double timeout_s, void** handle) {
  if (world > kMaxWorld || channel < 0 || channel >= kMaxChannels || blocks < 1 || blocks > kMaxBlocks) {
    pdl_set_error("bucket: bad world/channel/block count (%d, %d, %d)", world, channel, blocks);
    return -1;
  }
  if (algo == NVLS && (!mc_in || !mc_out)) { pdl_set_error("bucket: nvls needs multicast pointers"); return -1; }
  if (algo == PUSH && inbox_row4 < n4) { pdl_set_error("bucket: push inbox rows shorter than the bucket"); return -1; }
  BucketHandle* h = new BucketHandle();
  for (int q = 0; q < world; ++q) {
    h->p.in[q] = in[q]; h->p.out[q] = out[q]; h->p.pad[q] = pad ? pad[q] : nullptr; h->p.ring[q] = ring ? ring[q] : nullptr;
  }
  h->p.mc_in = mc_in; h->p.mc_out = mc_out; h->p.state = static_cast<CommState*>(state);
  h->p.n4 = n4; h->p.row4 = inbox_row4; h->p.rank = rank; h->p.world = world; h->p.channel = channel;
  h->p.scale = static_cast<float>(scale); h->p.timeout_ns = static_cast<unsigned long long>(timeout_s * 1e9);
  h->algo = algo; h->blocks = blocks;
  PDL_CUDA_CHECK(cudaEventCreateWithFlags(&h->ready, cudaEventDisableTiming));
  PDL_CUDA_CHECK(cudaEventCreateWithFlags(&h->done, cudaEventDisableTiming));
  *handle = h;
  return 0;
}

PDL_API int pdl_bucket_destroy(void* handle) {
  BucketHandle* h = static_cast<BucketHandle*>(handle);
  if (!h) return 0;
  cudaEventDestroy(h->ready);
  cudaEventDestroy(h->done);
  delete h;
  return 0;
}

static int bucket_launch(const ARParams& p, int algo, int blocks, cudaStream_t stream) {
  const bool sgd = p.param != nullptr;
  DISPATCH_ALGO(allreduce_kernel, p, blocks, false)
}

PDL_API int pdl_bucket_reduce(void* handle, float* param, float* mom, void* hyper, cudaStream_t compute, cudaStream_t comm) {
  BucketHandle* h = static_cast<BucketHandle*>(handle);
  ARParams p = h->p;
  p.param = param; p.mom = mom; p.hyper = static_cast<SgdHyper*>(hyper);
  PDL_CUDA_CHECK(cudaEventRecord(h->ready, compute));
  PDL_CUDA_CHECK(cudaStreamWaitEvent(comm, h->ready, 0));
  if (int rc = bucket_launch(p, h->algo, h->blocks, comm)) return rc;
  PDL_CUDA_CHECK(cudaEventRecord(h->done, comm));
  return 0;
}

PDL_API int pdl_bucket_wait(void* handle, cudaStream_t stream) {
  BucketHandle* h = static_cast<BucketHandle*>(handle);
  PDL_CUDA_CHECK(cudaStreamWaitEvent(stream, h->done, 0));
  return 0;
}

// Test harness: run `world` virtual ranks on ONE GPU in a single cooperative launch.
// Pointer arrays are [world][world]: entry [v][q] = what virtual rank v uses for peer q.
PDL_API int pdl_allreduce_emulated(int algo, int world, long long n4, const float* const* in, float* const* out, float* const* ring,
                                   uint32_t* const* pad, float* const* param, float* const* mom, void* const* hyper, void* const* state,
                                   int blocks, cudaStream_t stream) {
  if (algo == NVLS) { pdl_set_error("nvls cannot be emulated on one GPU"); return -1; }
  EmuParams ep{};
  ep.blocks_per_rank = blocks;
  bool sgd = param && param[0];
  for (int v = 0; v < world; ++v) {
    ARParams& p = ep.r[v];
    for (int q = 0; q < world; ++q) { p.in[q] = in[q]; p.out[q] = out[q]; p.ring[q] = ring ? ring[q] : nullptr; p.pad[q] = pad[q]; }
    p.param = sgd ? param[v] : nullptr; p.mom = (sgd && mom) ? mom[v] : nullptr;
    p.hyper = hyper ? static_cast<SgdHyper*>(hyper[v]) : nullptr; p.state = static_cast<CommState*>(state[v]);
    p.n4 = n4; p.row4 = n4; p.rank = v; p.world = world; p.channel = 0; p.scale = 1.0f / world; p.timeout_ns = 5ull * 1000000000ull;
  }
  DISPATCH_ALGO(allreduce_emulated_kernel, ep, blocks * world, true)
}

PDL_API int pdl_pull_launch(int rank, int world, int channel, int root, long long n16, const void* const* src, void* dst, uint32_t* const* pad,
                            void* state, int blocks, double timeout_s, cudaStream_t stream) {
  CopyParams c{};
  for (int q = 0; q < world; ++q) { c.src[q] = static_cast<const float4*>(src[q]); c.pad[q] = pad[q]; }
  c.dst = static_cast<float4*>(dst); c.state = static_cast<CommState*>(state);
  c.n16 = n16; c.rank = rank; c.world = world; c.channel = channel; c.root = root;
  c.timeout_ns = static_cast<unsigned long long>(timeout_s * 1e9);
  if (blocks < 1) blocks = 1;
  if (blocks > kMaxBlocks) blocks = kMaxBlocks;
  pull_kernel<<<blocks, kThreads, 0, stream>>>(c);
  PDL_LAUNCH_CHECK();
  return 0;
}
