#!/usr/bin/env bash
# compute-sanitizer over the single-GPU kernel tests ("Race/hang" tier).
#
#   gpurun --timeout 1500 -- 'bash scripts/sanitize.sh memcheck'      (or racecheck / synccheck / initcheck)
#
# One tool per call (each multiplies the run time by 10-50x), never on a multi-rank command, and only after the same
# tests have passed without it.  The emulated multi-rank allreduce tests are excluded from racecheck: their ranks are
# cooperating CTAs that communicate through global-memory flags on purpose, which is outside racecheck's model
# (it tracks shared memory).  Summaries go to gpurun_out/sanitize_<tool>.log; copy what should be kept to profiles/.
set -u
tool="${1:-memcheck}"
out="gpurun_out/sanitize_${tool}.log"
mkdir -p gpurun_out
tests=(
  tests/test_gpu_ops.py
  tests/test_gpu_tc.py
  tests/test_gpu_tc_strided.py
  tests/test_gpu_linear.py
  tests/test_metrics.py
  tests/test_zz_no_alloc.py
)
select=""
if [ "$tool" = "racecheck" ] || [ "$tool" = "synccheck" ]; then
  select="not emulated"
fi
# plain run first: a failing test must not be reported as a sanitizer finding
if ! python -m pytest "${tests[@]}" -m gpu -x -q ${select:+-k "$select"} > "gpurun_out/sanitize_plain.log" 2>&1; then
  echo "plain run failed; see gpurun_out/sanitize_plain.log" | tee "$out"
  exit 1
fi
PDL_NO_PDL=1 timeout 1400 compute-sanitizer --tool "$tool" --error-exitcode 9 --print-limit 20 \
  python -m pytest "${tests[@]}" -m gpu -x -q ${select:+-k "$select"} > "$out" 2>&1
rc=$?
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" "$out" | tail -5
exit $rc
