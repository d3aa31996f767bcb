#!/usr/bin/env bash
# compute-sanitizer over the single-GPU kernel tests ("Race/hang" tier).
#
#   gpurun --timeout 900 -- 'bash scripts/sanitize.sh memcheck'      (or racecheck / synccheck / initcheck)
#
# One tool per call (each multiplies the run time by 2-100x), never on a multi-rank command, and only after the same
# tests have passed without it in the same call.  The selection covers every kernel family once -- tcgen05 forward with the
# fused ReLU+pool epilogue, dgrad with the pool-backward epilogue, MN-major wgrad + its deterministic finish, the strided /
# stem variants, the fused CUDA-core ConvNet kernels (through the eager fused blocks), BatchNorm fused with ReLU/residual,
# pooling, the emulated multi-rank allreduce incl. the push variant (memcheck / initcheck only: its ranks are cooperating
# CTAs that talk through global-memory flags on purpose, outside racecheck's shared-memory model), SGD, Linear, metrics and
# the allocation-free wgrad.  racecheck instruments every shared-memory access and is the slowest tool by far: it gets the
# tensor-core and fused kernels only.  Summaries go to gpurun_out/sanitize_<tool>.log; copy what should be kept to profiles/.
set -u
tool="${1:-memcheck}"
out="gpurun_out/sanitize_${tool}.log"
mkdir -p gpurun_out
files=(tests/test_gpu_tc.py tests/test_gpu_ops.py tests/test_gpu_tc_strided.py tests/test_gpu_linear.py tests/test_metrics.py tests/test_zz_no_alloc.py)
wide="fused_relu_pool or dgrad_with_fused_pool or (tc_wgrad_matches and (case0 or case1 or case2 or case3)) or eager_fused_blocks \
or (tc_forward_matches and (case0 or case1 or case3 or case7)) or (tc_dgrad_matches and (case0 or case3)) or (forward_stride2 and case0) \
or (dgrad_stride2 and case0) or (wgrad_stride2 and case0) or (stem_conv and case0) or batchnorm_fused or maxpool_fwd_bwd or avgpool \
or add_relu or multi_tensor_sgd or relu_linear or linear_rows or running_metrics or noalloc or conv_epilogue_batchnorm"
case "$tool" in
  racecheck) select="fused_relu_pool or dgrad_with_fused_pool or (tc_wgrad_matches and (case0 or case1)) or eager_fused_blocks or (batchnorm_fused and shape0)" ;;
  synccheck) select="$wide" ;;
  *) select="$wide or fused_sgd_emulated" ;;
esac
# plain run first: a failing test must not be reported as a sanitizer finding
if ! python -m pytest "${files[@]}" -m gpu -x -q -k "$select" > "gpurun_out/sanitize_plain_${tool}.log" 2>&1; then
  echo "plain run failed; see gpurun_out/sanitize_plain_${tool}.log" | tee "$out"
  tail -5 "gpurun_out/sanitize_plain_${tool}.log"
  exit 1
fi
tail -1 "gpurun_out/sanitize_plain_${tool}.log"
PDL_NO_PDL=1 timeout "${SANITIZE_TIMEOUT:-700}" compute-sanitizer --tool "$tool" --error-exitcode 9 --print-limit 20 \
  python -m pytest "${files[@]}" -m gpu -x -q -k "$select" > "$out" 2>&1
rc=$?
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|Error:|Hazard" "$out" | tail -12
echo "sanitizer exit code: $rc" | tee -a "$out"
exit $rc
