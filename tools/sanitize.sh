#!/bin/bash
# Race detection / memory checking.
#   tools/sanitize.sh host          ThreadSanitizer build of the C++ runtime + in-process multi-rank driver (CPU box)
#   tools/sanitize.sh host-asan     the same driver under AddressSanitizer + UndefinedBehaviorSanitizer
#   tools/sanitize.sh gpu [tool]    compute-sanitizer (memcheck | racecheck | synccheck | initcheck) over the kernel
#                                   smoke run, one GPU (run under gpurun); report -> gpurun_out/sanitizer_<tool>.log
# Cross-GPU flag protocols cannot be seen by racecheck; they are covered by tests/test_multigpu.py (bit-equality
# across ranks over back-to-back launches with wrapped parity and skewed ranks).
set -u
cd "$(dirname "$0")/.."
mode=${1:-host}
if [ "$mode" = host ]; then
  mkdir -p build/san
  g++ -std=c++17 -g -O1 -fsanitize=thread -pthread csrc/runtime/*.cc csrc/runtime_tests/tsan_driver.cc -o build/san/tsan_driver || exit 2
  TFDL_RING_THREAD_BYTES=65536 TSAN_OPTIONS="halt_on_error=1 exitcode=66" timeout 600 ./build/san/tsan_driver "${2:-4}" "${3:-24600}" "${4:-10}"
  exit $?
fi
if [ "$mode" = host-asan ]; then
  mkdir -p build/san
  g++ -std=c++17 -g -O1 -fsanitize=address,undefined -fno-omit-frame-pointer -pthread csrc/runtime/*.cc csrc/runtime_tests/tsan_driver.cc -o build/san/asan_driver || exit 2
  TFDL_RING_THREAD_BYTES=65536 ASAN_OPTIONS=detect_leaks=0 UBSAN_OPTIONS=halt_on_error=1:print_stacktrace=1 timeout 600 ./build/san/asan_driver "${2:-4}" "${3:-24700}" "${4:-6}"
  exit $?
fi
tool=${2:-memcheck}
mkdir -p gpurun_out
export TFDL_USE_GRAPH=0 B=16 ITERS=1
python tools/prof_kernels.py > gpurun_out/sanitizer_plain.log 2>&1 || { echo "plain run failed"; tail -5 gpurun_out/sanitizer_plain.log; exit 1; }
timeout 900 compute-sanitizer --tool "$tool" --error-exitcode 66 --print-limit 20 python tools/prof_kernels.py > "gpurun_out/sanitizer_${tool}.log" 2>&1
rc=$?
tail -8 "gpurun_out/sanitizer_${tool}.log"
echo "compute-sanitizer $tool rc=$rc"
exit $rc
