"""ctypes signatures of libpdl_cuda.so (mirrors include/pdl_b200.h)."""
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_longlong, c_uint64, c_void_p

P = c_void_p
I = c_int
LL = c_longlong
F = c_float


def declare(L):
    def sig(name, args, res=c_int):
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = res

    sig("pdl_last_error", [], c_char_p)
    sig("pdl_cuda_device_count", [])
    sig("pdl_launch_count", [], LL)
    # symmetric memory
    sig("pdl_symm_multicast_supported", [I])
    sig("pdl_symm_begin", [I, I, I, c_uint64, c_char_p, I, POINTER(P)])
    sig("pdl_symm_exchange", [P])
    sig("pdl_symm_multicast_ready", [P])
    sig("pdl_symm_bind_multicast", [P, I])
    sig("pdl_symm_bytes", [P], c_uint64)
    sig("pdl_symm_ptr", [P, I], P)
    sig("pdl_symm_mc_ptr", [P], P)
    sig("pdl_symm_dev_ptrs", [P], P)
    sig("pdl_symm_destroy", [P])
    # collectives
    sig("pdl_comm_state_bytes", [])
    sig("pdl_comm_state_error_offset", [])
    sig("pdl_comm_state_host_error_offset", [])
    sig("pdl_signal_pad_bytes", [])
    sig("pdl_comm_max_blocks", [])
    sig("pdl_comm_max_channels", [])
    sig("pdl_sgd_hyper_bytes", [])
    sig("pdl_allreduce_blocks", [LL, I, I])
    sig("pdl_allreduce_push_inbox_bytes", [LL, I], LL)
    sig("pdl_allreduce_launch", [I, I, I, I, LL, P, P, P, P, P, P, P, P, P, P, I, c_double, c_double, P])
    sig("pdl_allreduce_launch_ex", [I, I, I, I, LL, LL, P, P, P, P, P, P, P, P, P, P, I, c_double, c_double, P])
    sig("pdl_bucket_create", [I, I, I, I, LL, LL, P, P, P, P, P, P, P, I, c_double, c_double, POINTER(P)])
    sig("pdl_bucket_reduce", [P, P, P, P, P, P])
    sig("pdl_bucket_wait", [P, P])
    sig("pdl_bucket_destroy", [P])
    sig("pdl_allreduce_emulated", [I, I, LL, P, P, P, P, P, P, P, P, I, P])
    sig("pdl_pull_launch", [I, I, I, I, LL, P, P, P, P, I, c_double, P])
    # optimizer
    sig("pdl_sgd_chunk_elems", [])
    sig("pdl_multi_tensor_sgd", [P, P, I, I, F, F, F, F, I, I, P])
    # convolution (CUDA-core fallback / oracle)
    sig("pdl_conv2d_fwd_simt", [P, P, P, P, I, I, I, I, I, I, I, I, I, I, P])
    sig("pdl_conv2d_dgrad_simt", [P, P, P, I, I, I, I, I, I, I, I, I, P])
    sig("pdl_conv2d_wgrad_scratch_bytes", [I, I, I, I, I, I, I, I, I], LL)
    sig("pdl_conv2d_wgrad_simt", [P, P, P, P, P, I, I, I, I, I, I, I, I, I, P])
    # convolution (tcgen05 / TMA tensor-core path)
    sig("pdl_tc_conv_supported", [I, I, I, I, I, I, I, I, I])
    sig("pdl_tc_conv_max_ctas", [I])
    sig("pdl_tc_pack_weights", [P, P, P, I, I, I, P])
    sig("pdl_tc_pack_max_taps", [])
    sig("pdl_tc_pack_weights_multi", [P, P, I, I, P])
    sig("pdl_tc_conv2d_fwd", [P, P, P, P, P, I, I, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_tc_conv2d_fwd_strided", [P, P, P, P, I, I, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_tc_conv2d_dgrad", [P, P, P, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_tc_wgrad_supported", [I, I, I, I, I, I, I, I, I])
    sig("pdl_tc_wgrad_max_ctas", [I])
    sig("pdl_wgrad_finish_comm_inbox_floats", [I, I, I, I, I], LL)
    sig("pdl_wgrad_finish_comm_blocks", [I, I, I, I])
    sig("pdl_tc_conv2d_wgrad_finish_comm", [P, P, I, I, I, I, I, I, I, I, I, P, P, P, P, P, I, I, I, P, P, P, c_double, c_double, P])
    sig("pdl_tc_conv2d_wgrad_gemm", [P, P, P, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_tc_conv2d_wgrad_finish", [P, P, P, I, P, I, I, I, I, I, I, I, I, P])
    sig("pdl_tc_wgrad_scratch_bytes", [I, I, I, I, I, I, I, I], LL)
    sig("pdl_tc_conv2d_dgrad_strided", [P, P, P, I, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_tc_wgrad_scratch_bytes_strided", [I, I, I, I, I, I, I, I, I], LL)
    sig("pdl_tc_conv2d_wgrad_strided", [P, P, P, P, P, I, P, I, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_tc_conv2d_wgrad", [P, P, P, P, P, I, P, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_tc_conv2d_dgrad_add", [P, P, P, P, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_tc_conv_bn_splits", [I, I, I, I, I, I, I, I, I])
    sig("pdl_tc_conv2d_fwd_bnstats", [P, P, P, P, I, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_tc_stem_bn_splits", [I, I, I, I, I, I, I, I, I])
    sig("pdl_tc_stem_fwd_bnstats", [P, P, P, P, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_bn_fwd_train_prestats", [P, P, P, P, P, P, P, F, F, P, P, P, P, I, LL, I, I, P])
    sig("pdl_tc_dgrad_poolbwd_tiles", [I, I, I])
    sig("pdl_tc_conv2d_dgrad_poolbwd", [P, P, P, P, P, P, I, I, I, I, I, I, I, I, I, P, P])
    # small-Cin ("stem") convolutions on the tensor cores
    sig("pdl_tc_stem_supported", [I, I, I, I, I, I, I, I])
    sig("pdl_tc_stem_pad_floats", [I, I, I, I, I, I, I], LL)
    sig("pdl_tc_stem_prepare", [P, LL, LL, LL, LL, P, I, I, I, I, I, I, I, I, P])
    sig("pdl_tc_stem_pack_weights", [P, P, I, I, I, I, P])
    sig("pdl_tc_stem_fwd", [P, P, P, P, I, I, I, I, I, I, I, I, I, P, P])
    sig("pdl_tc_stem_wgrad_supported", [I, I, I, I, I, I, I, I, I])
    sig("pdl_tc_stem_wgrad_scratch_bytes", [I, I, I, I, I, I, I, I, I], LL)
    sig("pdl_tc_stem_wgrad", [P, P, P, P, I, I, I, I, I, I, I, I, I, P, P])
    # fused ConvNet kernels
    sig("pdl_convnet_pack", [P, P, P, P, P, P, P, P, P])
    sig("pdl_conv1_fused_fwd", [P, P, P, P, P, I, P])
    sig("pdl_fc_ce_fused", [P, P, P, P, P, P, P, P, I, P])
    sig("pdl_fc_bwd_splits", [I])
    sig("pdl_fc_bwd_scratch_floats", [I], LL)
    sig("pdl_fc_bwd_pool3_fused", [P, P, P, P, P, P, P, P, P, I, P])
    sig("pdl_fc_bwd_pool3_part", [I, P, P, P, P, P, P, P, P, P, I, P])
    sig("pdl_fc_ce_bwd_fused", [P, P, P, P, P, P, P, P, P, P, P, I, P])
    sig("pdl_bwd_blocks", [I])
    sig("pdl_pool2_bwd_blocks", [I, I, I, I])
    sig("pdl_pool2_bwd_dbias", [P, P, P, P, P, I, I, I, I, I, P])
    sig("pdl_conv1_wgrad_fused", [P, P, P, P, P, P, P, I, P])
    sig("pdl_conv1_tail_inbox_bytes", [I], LL)
    sig("pdl_conv1_wgrad_fused_comm", [P, P, P, P, P, P, P, I, P, P, P, P, P, P, P, I, I, P, P, P, c_double, c_double, P])
    # batch norm
    sig("pdl_bn_scratch_floats", [LL, I], LL)
    sig("pdl_bn_fwd_train", [P, P, P, P, P, P, P, F, F, P, P, P, P, LL, I, I, P])
    sig("pdl_bn_fwd_eval", [P, P, P, P, P, P, F, P, P, LL, I, I, P])
    sig("pdl_bn_bwd", [P, P, P, P, P, P, P, P, P, P, LL, I, I, P])
    sig("pdl_bn_bwd2", [P, P, P, P, P, P, P, P, P, P, P, LL, I, I, P])
    # element-wise / pooling / linear / loss
    sig("pdl_relu_fwd", [P, P, LL, P])
    sig("pdl_relu_bwd", [P, P, P, LL, P])
    sig("pdl_add_relu_fwd", [P, P, P, LL, P])
    sig("pdl_maxpool_fwd", [P, P, P, I, I, I, I, I, I, I, P])
    sig("pdl_maxpool_bwd", [P, P, P, P, I, I, I, I, I, I, I, I, P])
    sig("pdl_global_avgpool_fwd", [P, P, I, I, I, P])
    sig("pdl_global_avgpool_bwd", [P, P, I, I, I, P])
    sig("pdl_permute_nhwc_nchw", [P, P, I, I, I, I, P])
    sig("pdl_colsum_splits", [LL, I])
    sig("pdl_colsum_partial", [P, P, LL, I, P])
    sig("pdl_linear_fwd", [P, P, P, P, I, I, I, P])
    sig("pdl_linear_dgrad", [P, P, P, I, I, I, P])
    sig("pdl_linear_wgrad", [P, P, P, P, I, I, I, P])
    sig("pdl_cross_entropy", [P, P, I, I, P, P, P, P, P])
    sig("pdl_step_metrics", [P, P, P, I, I, P, P])
    sig("pdl_copy_pair", [P, P, LL, P, P, LL, P])
    for name, args, res in _OPTIONAL:
        if hasattr(L, name):
            sig(name, args, res)


# entry points added by later kernels register here: (name, argtypes, restype)
_OPTIONAL = []


def optional(name, args, res=c_int):
    _OPTIONAL.append((name, args, res))
