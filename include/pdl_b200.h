/* C ABI of libpdl_cuda.so -- sm_100a kernels and the symmetric-memory runtime.
 *
 * Every launcher enqueues on the given stream and returns 0 or an error code (message via
 * pdl_last_error()).  Activations are NHWC fp32; convolution weights are [Cout][Cin][KH][KW].
 * "reference:" cites the line of Jackxiini/Pytorch-distributed-learning whose behaviour the
 * entry point provides (the reference itself ships no native code; it reaches cuDNN/cuBLAS/NCCL
 * through PyTorch).
 */
#ifndef PDL_B200_H_
#define PDL_B200_H_

#include <cuda_runtime.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

const char* pdl_last_error(void);
int pdl_cuda_device_count(void);
long long pdl_launch_count(void); /* launcher calls made by this process so far (each enqueues >= 1 kernel) */

/* ---- symmetric memory: same-size VMM allocation on every rank, peer-mapped (+ NVSwitch multicast) ----
 * reference: the NCCL communicator behind init_process_group(backend='nccl') (launch_dist.py:49).
 * Phases are separated by caller-side barriers: begin -> exchange -> bind_multicast. */
int pdl_symm_multicast_supported(int device);
int pdl_symm_begin(int rank, int world, int device, uint64_t bytes, const char* uds_prefix, int want_multicast, void** out);
int pdl_symm_exchange(void* seg);
int pdl_symm_multicast_ready(void* seg);
int pdl_symm_bind_multicast(void* seg, int all_ok);
uint64_t pdl_symm_bytes(void* seg);
void* pdl_symm_ptr(void* seg, int rank); /* address through which THIS process reaches rank's copy */
void* pdl_symm_mc_ptr(void* seg);        /* multicast alias or NULL */
void** pdl_symm_dev_ptrs(void* seg);
int pdl_symm_destroy(void* seg);

/* ---- gradient-bucket allreduce fused with 1/W averaging and SGD (reference: loss.backward() ->
 *      DDP bucket allreduce -> optimizer.step(), launch_dist.py:91-92) ----
 * algo: 0 one-shot, 1 two-shot, 2 ring, 3 NVLS multimem, 4 push (stores into the peers' inboxes, one flag round).
 * in/out/ring/pad: `world` peer-mapped pointers; `ring` is the ring's staging buffer or, for push, the bucket's own inbox
 * of pdl_allreduce_push_inbox_bytes(n4, world) bytes per rank.
 * param/mom/hyper non-NULL => the SGD update (weight decay, momentum, nesterov) runs in the same kernel. */
int pdl_comm_state_bytes(void);
/* byte offsets inside the communicator state: the sticky error word (1 = a peer wait timed out; every later launch on the
 * communicator then leaves gradients and parameters untouched), and an optional device-visible pinned host word that the
 * failing kernel also sets so the host can notice with a plain load */
int pdl_comm_state_error_offset(void);
int pdl_comm_state_host_error_offset(void);
int pdl_signal_pad_bytes(void);
int pdl_comm_max_blocks(void);
int pdl_comm_max_channels(void);
int pdl_sgd_hyper_bytes(void);
int pdl_allreduce_blocks(long long n4, int world, int algo);
long long pdl_allreduce_push_inbox_bytes(long long n4, int world);
int pdl_allreduce_launch(int algo, int rank, int world, int channel, long long n4, const float* const* in, float* const* out,
                         float* const* ring, uint32_t* const* pad, const float* mc_in, float* mc_out, float* param, float* mom, void* hyper,
                         void* state, int blocks, double scale, double timeout_s, cudaStream_t stream);
/* the same with an explicit distance between inbox rows (push variant; an inbox shared by launches of different lengths) */
int pdl_allreduce_launch_ex(int algo, int rank, int world, int channel, long long n4, long long inbox_row4, const float* const* in,
                            float* const* out, float* const* ring, uint32_t* const* pad, const float* mc_in, float* mc_out, float* param,
                            float* mom, void* hyper, void* state, int blocks, double scale, double timeout_s, cudaStream_t stream);
/* The reducer's launch path as one native call per bucket (reference: Reducer::mark_bucket_ready -> allreduce on the process
 * group's stream).  A handle fixes peer pointers, variant and CTA count; pdl_bucket_reduce records "gradients ready" on the
 * compute stream, makes the communication stream wait for it, launches the fused kernel there and records "done";
 * pdl_bucket_wait orders a stream behind "done".  Capturable. */
int pdl_bucket_create(int algo, int rank, int world, int channel, long long n4, long long inbox_row4, const float* const* in,
                      float* const* out, float* const* ring, uint32_t* const* pad, const float* mc_in, float* mc_out, void* state, int blocks,
                      double scale, double timeout_s, void** handle);
int pdl_bucket_reduce(void* handle, float* param, float* mom, void* hyper, cudaStream_t compute, cudaStream_t comm);
int pdl_bucket_wait(void* handle, cudaStream_t stream);
int pdl_bucket_destroy(void* handle);
/* all `world` ranks as one cooperative launch on one GPU (tests) */
int pdl_allreduce_emulated(int algo, int world, long long n4, const float* const* in, float* const* out, float* const* ring,
                           uint32_t* const* pad, float* const* param, float* const* mom, void* const* hyper, void* const* state, int blocks,
                           cudaStream_t stream);
/* broadcast (root >= 0) / allgather (root < 0) by pulling from peers' staging buffers (reference: DDP
 * constructor broadcast + per-forward buffer broadcast, launch_dist.py:61, example_launch.py:29) */
int pdl_pull_launch(int rank, int world, int channel, int root, long long n16, const void* const* src, void* dst, uint32_t* const* pad,
                    void* state, int blocks, double timeout_s, cudaStream_t stream);

/* ---- stand-alone SGD (reference: torch.optim.SGD at launch_dist.py:59, example_launch.py:64-70) ---- */
int pdl_sgd_chunk_elems(void);
int pdl_multi_tensor_sgd(const void* table, const int* chunk_prefix, int num_tensors, int total_chunks, float lr, float momentum,
                         float dampening, float weight_decay, int nesterov, int first_step, cudaStream_t stream);

/* ---- convolution, CUDA cores, strict fp32, any shape (reference: nn.Conv2d at launch_dist.py:15-25) ---- */
int pdl_conv2d_fwd_simt(const float* x, const float* w, const float* bias, float* y, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                        int stride, int pad, int relu, cudaStream_t stream);
int pdl_conv2d_dgrad_simt(const float* dy, const float* w, float* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad,
                          cudaStream_t stream);
long long pdl_conv2d_wgrad_scratch_bytes(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad);
int pdl_conv2d_wgrad_simt(const float* x, const float* dy, float* dw, float* db, float* scratch, int N, int H, int W, int Cin, int Cout, int KH,
                          int KW, int stride, int pad, cudaStream_t stream);

/* Allocation-free twin of pdl_conv2d_wgrad_simt (no counterpart in the reference; it exists for embedding
 * hosts, not for parity).  Who needs it: processes whose device memory is already fully reserved by another
 * allocator (a framework's caching pool, a serving runtime's arena), where a fresh cudaMalloc fails, and
 * latency-critical loops that must stay away from cudaMalloc/cudaFree because those may synchronise the
 * device.  The split-K partial sums go to an 8 MiB array built into the library (resident on every GPU that
 * loads it, for the life of the process); pdl_conv2d_wgrad_noalloc_max_bytes() is the limit and larger shapes
 * are refused with an error.  The array is shared: calls that can overlap on the GPU race, keep them
 * stream-ordered.  The first call on each device resolves the array's device address; make it before any
 * stream capture starts.  Safe to call from several host threads. */
long long pdl_conv2d_wgrad_noalloc_max_bytes(void);
int pdl_conv2d_wgrad_simt_noalloc(const float* x, const float* dy, float* dw, float* db, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                                  int stride, int pad, cudaStream_t stream);

/* ---- convolution, tcgen05 tensor cores (TF32 in, fp32 accumulate in TMEM), TMA-fed (same reference lines) ---- */
int pdl_tc_conv_supported(int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad, int pool);
int pdl_tc_conv_max_ctas(int cap); /* bound the CTA count of the forward / dgrad launches that follow (0 = none); returns the old bound */
int pdl_tc_pack_weights(const float* w, float* packed_fwd, float* packed_bwd, int Cout, int Cin, int taps, cudaStream_t stream);
/* every convolution weight of a model in one launch (shared-memory tile transpose, both operand layouts): table
 * [num][6] int64 {w, fwd, bwd, Cout, Cin, taps} (device; channels % 32 == 0, taps <= pdl_tc_pack_max_taps()),
 * prefix [num+1] int32 first 32x32-channel tile of each tensor */
int pdl_tc_pack_max_taps(void);
int pdl_tc_pack_weights_multi(const void* table, const int* prefix, int num, int total_tiles, cudaStream_t stream);
/* pool_stride 0: y = conv+bias(+relu); 1|2: fused ReLU + 2x2 max-pool, y/idx are the pooled output and argmax */
int pdl_tc_conv2d_fwd(const float* x, const float* packed_w, const float* bias, float* y, uint8_t* idx, int N, int H, int W, int Cin, int Cout,
                      int KH, int KW, int pad, int relu, int pool_stride, int* error_flag, cudaStream_t stream);
int pdl_tc_conv2d_fwd_strided(const float* x, const float* packed_w, const float* bias, float* y, int N, int H, int W, int Cin, int Cout, int KH,
                              int KW, int stride, int pad, int relu, int* error_flag, cudaStream_t stream); /* stride 1 or 2 */
int pdl_tc_conv2d_dgrad(const float* dy, const float* packed_w_bwd, float* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad,
                        int* error_flag, cudaStream_t stream);
int pdl_tc_wgrad_supported(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad);
/* bound the CTA count of the weight-gradient launches that follow (0 = no bound; returns the previous bound): for small
 * weight gradients that run next to other kernels, see tc_wgrad.cu */
int pdl_tc_wgrad_max_ctas(int cap);
long long pdl_tc_wgrad_scratch_bytes(int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad);
int pdl_tc_conv2d_wgrad(const float* x, const float* dy, float* dw, float* db, const float* db_partial, int db_blocks, float* scratch, int N,
                        int H, int W, int Cin, int Cout, int KH, int KW, int pad, int* error_flag, cudaStream_t stream);
/* the two passes of pdl_tc_conv2d_wgrad (split-K GEMM into `scratch`; deterministic reduction of the partials + bias gradient)
 * as separate launches, for callers that schedule them on different streams */
int pdl_tc_conv2d_wgrad_gemm(const float* x, const float* dy, float* scratch, int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad,
                             int* error_flag, cudaStream_t stream);
int pdl_tc_conv2d_wgrad_finish(float* dw, float* db, const float* db_partial, int db_blocks, const float* scratch, int N, int H, int W, int Cin,
                               int Cout, int KH, int KW, int pad, cudaStream_t stream);
/* second pass of pdl_tc_conv2d_wgrad_gemm FUSED with the cross-GPU exchange of the layer's gradient and its SGD update (reference:
 * gradient -> bucket allreduce -> optimizer.step(), launch_dist.py:91-92, for one layer in one kernel): every CTA reduces its
 * split-K partials, pushes the result into the peers' inboxes (symmetric memory, [2][world][Cout + Cout*Cin*KH*KW] floats),
 * raises one flag per peer, sums the rows it received in rank order and updates the parameters.  raw/avg/param/mom: the
 * layer's [db | dw] slice of the gradient / averaged-gradient / parameter / momentum arenas. */
long long pdl_wgrad_finish_comm_inbox_floats(int world, int Cin, int Cout, int KH, int KW);
int pdl_wgrad_finish_comm_blocks(int Cin, int Cout, int KH, int KW);
int pdl_tc_conv2d_wgrad_finish_comm(const float* scratch, const float* db_partial, int db_blocks, int N, int H, int W, int Cin, int Cout, int KH,
                                    int KW, int pad, float* raw, float* avg, float* param, float* mom, void* hyper, int rank, int world,
                                    int channel, float* const* inbox, uint32_t* const* sigpad, void* state, double scale, double timeout_s,
                                    cudaStream_t stream);
int pdl_tc_conv2d_dgrad_strided(const float* dy, const float* packed_w_bwd, float* dx, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                                int stride, int pad, int* error_flag, cudaStream_t stream); /* stride 2 = four parity-class launches */
long long pdl_tc_wgrad_scratch_bytes_strided(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad);
int pdl_tc_conv2d_wgrad_strided(const float* x, const float* dy, float* dw, float* db, const float* db_partial, int db_blocks, float* scratch,
                                int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad, int* error_flag,
                                cudaStream_t stream); /* stride 1 or 2 */
/* dx = dgrad(dy, w) + add_src, stride 1: the skip connection's gradient of a residual block is added in the epilogue of
 * the block's first convolution (returns -5 when the TMA-store epilogue is switched off) */
int pdl_tc_conv2d_dgrad_add(const float* dy, const float* packed_w_bwd, const float* add_src, float* dx, int N, int H, int W, int Cin, int Cout,
                            int KH, int KW, int pad, int* error_flag, cudaStream_t stream);
/* Bias-free forward that also leaves the BatchNorm statistics of its output in bn_partial ([splits][Cout][3] floats:
 * count, mean, M2 per CTA group; splits = pdl_tc_conv_bn_splits(...), 0 = shape not covered): the epilogue reduces
 * the staged output tile, so the following BatchNorm (pdl_bn_fwd_train_prestats) does not re-read the tensor. */
int pdl_tc_conv_bn_splits(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad);
int pdl_tc_conv2d_fwd_bnstats(const float* x, const float* packed_w, float* y, float* bn_partial, int N, int H, int W, int Cin, int Cout, int KH,
                              int KW, int stride, int pad, int* error_flag, cudaStream_t stream);
/* dgrad of a stride-1 conv whose input is ReLU + 2x2 max-pool (stride 1|2) of the layer below, pool/ReLU backward fused
 * into the epilogue: writes dz[N][(H-1)*ps+2][(W-1)*ps+2][Cin] (TF32-rounded) and dbpart[tiles][Cin] column sums */
int pdl_tc_dgrad_poolbwd_tiles(int N, int H, int W);
int pdl_tc_conv2d_dgrad_poolbwd(const float* dy, const float* packed_w_bwd, const uint8_t* idx, const float* pooled, float* dz, float* dbpart,
                                int N, int H, int W, int Cin, int Cout, int KH, int KW, int pad, int pool_stride, int* error_flag,
                                cudaStream_t stream);
void pdl_tc_set_trace(long long* buf); /* profiling aid: clock64() milestones of CTA 0 */

/* ---- small-Cin ("stem") convolutions on the tensor cores: Cin <= 4, KW <= 8, stride 1|2 (reference: the 7x7/2
 *      3-channel conv1 of torchvision resnet18, example_launch.py:26).  The input is first copied into a zero-padded
 *      NHWC4 image (pdl_tc_stem_prepare) in which each filter row is a contiguous 32-float K chunk. ---- */
int pdl_tc_stem_supported(int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad);
long long pdl_tc_stem_pad_floats(int N, int H, int W, int KH, int KW, int stride, int pad);
int pdl_tc_stem_prepare(const float* x, long long sn, long long sc, long long sh, long long sw, float* xp, int N, int Cin, int H, int W, int KH,
                        int KW, int stride, int pad, cudaStream_t stream);
int pdl_tc_stem_pack_weights(const float* w, float* packed, int Cout, int Cin, int KH, int KW, cudaStream_t stream);
int pdl_tc_stem_fwd(const float* xp, const float* packed_w, const float* bias, float* y, int N, int H, int W, int Cout, int KH, int KW, int stride,
                    int pad, int relu, int* error_flag, cudaStream_t stream);
int pdl_tc_stem_bn_splits(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad);
int pdl_tc_stem_fwd_bnstats(const float* xp, const float* packed_w, float* y, float* bn_partial, int N, int H, int W, int Cout, int KH, int KW,
                            int stride, int pad, int* error_flag, cudaStream_t stream); /* stem forward + BatchNorm statistics of y */
int pdl_tc_stem_wgrad_supported(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad);
long long pdl_tc_stem_wgrad_scratch_bytes(int N, int H, int W, int Cin, int Cout, int KH, int KW, int stride, int pad);
int pdl_tc_stem_wgrad(const float* xp, const float* dy, float* dw, float* scratch, int N, int H, int W, int Cin, int Cout, int KH, int KW,
                      int stride, int pad, int* error_flag, cudaStream_t stream);

/* ---- fused ConvNet kernels (reference: ConvNet.forward launch_dist.py:35-41 and its autograd) ---- */
int pdl_convnet_pack(const float* w2, const float* w3, const float* wf, float* p2f, float* p2b, float* p3f, float* p3b, float* wfp,
                     cudaStream_t stream);
int pdl_conv1_fused_fwd(const float* x, const float* w, const float* bias, float* a1, uint8_t* idx1, int N, cudaStream_t stream);
int pdl_fc_ce_fused(const float* a3, const float* wfp, const float* bf, const long long* labels, float* logits, float* row_loss, float* loss,
                    float* dlogits, int N, cudaStream_t stream);
int pdl_fc_bwd_splits(int N);
long long pdl_fc_bwd_scratch_floats(int N);
int pdl_fc_bwd_pool3_fused(const float* a3, const uint8_t* idx3, const float* dlogits, const float* wfp, float* gwf, float* gbf, float* wpart,
                           float* dz3, float* db3part, int N, cudaStream_t stream);
int pdl_bwd_blocks(int N);
int pdl_fc_ce_bwd_fused(const float* a3, const float* wfp, const float* bf, const long long* labels, float* logits, float* row_loss, float* loss,
                        float* dlogits, const uint8_t* idx3, float* dz3, float* db3part, int N, cudaStream_t stream); /* fc1+loss+dz3 in one kernel */
int pdl_fc_bwd_pool3_part(int part, const float* a3, const uint8_t* idx3, const float* dlogits, const float* wfp, float* gwf, float* gbf,
                          float* wpart, float* dz3, float* db3part, int N, cudaStream_t stream); /* 1: fc1 wgrad only, 2: dz3/db3part only */
int pdl_pool2_bwd_blocks(int N, int H, int W, int C);
int pdl_pool2_bwd_dbias(const float* da, const uint8_t* idx, const float* pooled, float* dz, float* dbpart, int N, int H, int W, int C,
                        int stride, cudaStream_t stream);
int pdl_conv1_wgrad_fused(const float* x, const float* da1, const uint8_t* idx1, const float* a1, float* gw1, float* gb1, float* partial, int N,
                          cudaStream_t stream);
/* The same kernel finishing with the cross-GPU exchange of its result and the SGD update (reference: the last gradient ->
 * DDP bucket allreduce -> optimizer.step(), launch_dist.py:91-92).  conv1's gradient is the last one backward produces, so its
 * reduction cannot overlap anything: the CTA that holds the finished 832 values pushes them into every peer's inbox
 * (symmetric memory: [2][world][832] floats behind a signal pad of its own), raises ONE flag per peer, sums the rows it
 * received in rank order and updates conv1's parameters.  world == 1: no exchange.  param_w NULL: average only. */
long long pdl_conv1_tail_inbox_bytes(int world);
int pdl_conv1_wgrad_fused_comm(const float* x, const float* da1, const uint8_t* idx1, const float* a1, float* gw1, float* gb1, float* partial,
                               int N, float* avg_w, float* avg_b, float* param_w, float* param_b, float* mom_w, float* mom_b, void* hyper, int rank,
                               int world, float* const* inbox, uint32_t* const* pad, void* state, double scale, double timeout_s,
                               cudaStream_t stream);

/* ---- BatchNorm2d with the block's ReLU / residual add fused in, NHWC rows [M = N*H*W][C]:
 *      y = act(bn(x) [+ residual]) (reference: torchvision resnet18 at example_launch.py:26).
 *      saved is [4][C] = mean, invstd, scale, shift; residual / dres / y may be NULL as documented in bn.cu ---- */
long long pdl_bn_scratch_floats(long long M, int C);
int pdl_bn_fwd_train(const float* x, const float* residual, const float* gamma, const float* beta, float* running_mean, float* running_var,
                     long long* num_batches_tracked, float momentum, float eps, float* y, float* saved, uint8_t* relu_mask, float* scratch,
                     long long M, int C, int relu, cudaStream_t stream); /* relu_mask: M*C/4 bytes, written for the residual+ReLU form */
int pdl_bn_fwd_train_prestats(const float* x, const float* residual, const float* gamma, const float* beta, float* running_mean,
                              float* running_var, long long* num_batches_tracked, float momentum, float eps, float* y, float* saved,
                              uint8_t* relu_mask, const float* partial, int splits, long long M, int C, int relu, cudaStream_t stream);
int pdl_bn_fwd_eval(const float* x, const float* residual, const float* gamma, const float* beta, const float* running_mean,
                    const float* running_var, float eps, float* y, float* scratch, long long M, int C, int relu, cudaStream_t stream);
/* dy2 (may be NULL): a second tensor added to dy on the fly -- the gradient of a block output that feeds both the next
 * block's first convolution and its skip connection arrives in two pieces; reading both here replaces autograd's add pass */
int pdl_bn_bwd2(const float* x, const float* dy, const float* dy2, const uint8_t* relu_mask, const float* gamma, const float* saved, float* dx,
                float* dres, float* dgamma, float* dbeta, float* scratch, long long M, int C, int relu, cudaStream_t stream);
int pdl_bn_bwd(const float* x, const float* dy, const uint8_t* relu_mask, const float* gamma, const float* saved, float* dx, float* dres,
               float* dgamma, float* dbeta, float* scratch, long long M, int C, int relu, cudaStream_t stream);

/* ---- element-wise / pooling / linear / loss (reference: nn.ReLU, nn.MaxPool2d, nn.Linear,
 *      nn.CrossEntropyLoss at launch_dist.py:13-32,58; ResNet add+ReLU / avgpool at example_launch.py:26) ---- */
int pdl_relu_fwd(const float* x, float* y, long long n, cudaStream_t stream);
int pdl_relu_bwd(const float* y, const float* dy, float* dx, long long n, cudaStream_t stream);
int pdl_add_relu_fwd(const float* a, const float* b, float* y, long long n, cudaStream_t stream);
int pdl_maxpool_fwd(const float* x, float* y, uint8_t* idx, int N, int H, int W, int C, int k, int stride, int pad, cudaStream_t stream);
int pdl_maxpool_bwd(const float* dy, const uint8_t* idx, const float* pooled, float* dx, int N, int H, int W, int C, int k, int stride, int pad,
                    int fuse_relu, cudaStream_t stream);
int pdl_global_avgpool_fwd(const float* x, float* y, int N, int HW, int C, cudaStream_t stream);
int pdl_global_avgpool_bwd(const float* dy, float* dx, int N, int HW, int C, cudaStream_t stream);
int pdl_permute_nhwc_nchw(const float* x, float* y, int N, int HW, int C, int to_nchw, cudaStream_t stream);
/* part[pdl_colsum_splits(M, C)][C]: column sums of NHWC rows x[M][C] per row split -- a convolution's bias gradient as partial
 * sums for the weight-gradient kernel's second pass (reference: the db reduction of nn.Conv2d, launch_dist.py:15-25) */
int pdl_colsum_splits(long long M, int C);
int pdl_colsum_partial(const float* x, float* part, long long M, int C, cudaStream_t stream);
int pdl_linear_fwd(const float* x, const float* w, const float* b, float* y, int M, int N, int K, cudaStream_t stream);
int pdl_linear_dgrad(const float* dy, const float* w, float* dx, int M, int N, int K, cudaStream_t stream);
int pdl_linear_wgrad(const float* x, const float* dy, float* dw, float* db, int M, int N, int K, cudaStream_t stream);
int pdl_cross_entropy(const float* logits, const long long* target, int N, int C, float* row_loss, float* loss, float* dlogits,
                      const float* grad_scale, cudaStream_t stream);
/* Running metrics on the device (reference example_launch.py:91-93: argmax/eq/sum/.item() + loss.item() every step):
 * acc[0] += *loss (skipped if NULL), acc[1] += rows whose arg-max equals the label, acc[2] += N.  No host sync. */
/* Two device-to-device copies in one launch (a graphed step's images + labels; reference launch_dist.py:82-83). */
int pdl_copy_pair(void* d0, const void* s0, long long n0, void* d1, const void* s1, long long n1, cudaStream_t stream);
int pdl_step_metrics(const float* logits, const long long* target, const float* loss, int N, int C, double* acc, cudaStream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* PDL_B200_H_ */
