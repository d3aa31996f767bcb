#!/usr/bin/env bash
# Host runtime (csrc/host: store server/client threads, TCP collectives, bucket assignment, reducer) under
# AddressSanitizer + UBSan, or ThreadSanitizer.  Runs on any box, no GPU:
#
#   bash scripts/sanitize_host.sh asan      # heap / stack / UB errors
#   bash scripts/sanitize_host.sh tsan      # data races between the store's server threads and its clients
#
# The sanitizer build goes to build/ (the in-tree libpdl_host.so is untouched); the package loads it through PDL_HOST_LIB
# and the sanitizer runtime is pre-loaded into the (uninstrumented) python.  Spawned worker processes inherit both.
set -eu
kind="${1:-asan}"
root="$(cd "$(dirname "$0")/.." && pwd)"
mkdir -p "$root/build"
case "$kind" in
  asan) flags="-fsanitize=address,undefined -fno-sanitize-recover=undefined"; rt="$(gcc -print-file-name=libasan.so)"
        export ASAN_OPTIONS="detect_leaks=0:abort_on_error=0:exitcode=66" UBSAN_OPTIONS="print_stacktrace=1" ;;
  tsan) flags="-fsanitize=thread"; rt="$(gcc -print-file-name=libtsan.so)"
        # thread leaks are expected: the failure tests end their workers with os._exit() while the store's threads run
        export TSAN_OPTIONS="exitcode=66:report_signal_unsafe=0:report_thread_leaks=0" ;;
  *) echo "usage: $0 asan|tsan"; exit 2 ;;
esac
lib="$root/build/libpdl_host_${kind}.so"
# shellcheck disable=SC2086
g++ -O1 -g -fno-omit-frame-pointer $flags -std=c++17 -fPIC -shared -pthread -o "$lib" "$root"/csrc/host/*.cpp
export PDL_HOST_LIB="$lib" LD_PRELOAD="$rt"
cd "$root"
python -m pytest tests/test_store.py tests/test_bucket.py tests/test_reducer_core.py tests/test_cpu_backend.py tests/test_rendezvous.py \
  -x -q -m "not gpu" -p no:cacheprovider 2>&1 | tee "build/sanitize_host_${kind}.log" | tail -15
if grep -qE "ERROR: AddressSanitizer|runtime error:|WARNING: ThreadSanitizer" "build/sanitize_host_${kind}.log"; then
  echo "sanitizer findings: see build/sanitize_host_${kind}.log"; exit 1
fi
echo "host runtime: no ${kind} findings"
