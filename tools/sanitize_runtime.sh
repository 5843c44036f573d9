#!/bin/bash
# Host-runtime sanitizer pass: builds csrc/runtime/*.cpp with AddressSanitizer + UndefinedBehaviorSanitizer and
# runs the CPU tests that drive it (Reducer state machine, bucket assignment, partitioner, schedule generator / simulator,
# DDP on gloo) against that build.  Needs no GPU.  Usage: tools/sanitize_runtime.sh [extra pytest args]
set -euo pipefail
ROOT="$(cd "$(dirname "$0")/.." && pwd)"
OUT="${TMPDIR:-/tmp}/dmp_b200_rt_asan"
mkdir -p "$OUT"
g++ -O1 -g -std=c++17 -shared -fPIC -fno-omit-frame-pointer -fsanitize=address,undefined -fno-sanitize-recover=undefined \
    -Wall -I "$ROOT/include" -o "$OUT/libdmp_b200_rt_asan.so" "$ROOT"/distributed_model_parallel_b200/csrc/runtime/*.cpp -lpthread
ASAN_LIB="$(g++ -print-file-name=libasan.so)"
cd "$ROOT"
# python itself is not instrumented: preload the runtime, ignore interpreter leaks, abort on the first real finding
LD_PRELOAD="$ASAN_LIB" ASAN_OPTIONS="detect_leaks=0:abort_on_error=1:halt_on_error=1" UBSAN_OPTIONS="print_stacktrace=1:halt_on_error=1" \
DMP_B200_RT_LIB="$OUT/libdmp_b200_rt_asan.so" \
    python -m pytest tests/test_reducer.py tests/test_pipeline_cpu.py -q -x -p no:cacheprovider -k "not pipeline_matches_full_model" "$@"
