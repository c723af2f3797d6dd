#!/bin/bash
# compute-sanitizer passes over the small-shape kernel tests: memcheck on the fused cfconv, the grouped TN
# launch, the CSR primitives and the static-scratch column sum; racecheck on the cfconv and Dense kernels that hand
# shared memory between the generic and the async proxy by hand.  Each pass is bounded; logs go to gpurun_out/.
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
CS=/usr/local/cuda/bin/compute-sanitizer
run() {  # name tool limit pytest-args...
  local name=$1 tool=$2 limit=$3; shift 3
  local log=gpurun_out/r3_sanitizer_${name}.txt
  echo "== $tool: pytest $* (limit ${limit}s)" > "$log"
  timeout "$limit" $CS --tool "$tool" --print-limit 20 python -m pytest "$@" -x -q -m gpu -p no:cacheprovider >> "$log" 2>&1
  echo "exit code: $?" >> "$log"
  grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed|exit code" "$log" | tail -5
}
run memcheck_cfconv memcheck 420 tests/test_gpu_cfconv.py -k "c2 or hub or empty or sink"
run memcheck_tn_dense memcheck 420 tests/test_gpu_tn_grouped.py tests/test_zz_no_alloc.py
run memcheck_primitives memcheck 420 tests/test_gpu_primitives.py -k "pooling or gather or shift or segment"
run racecheck_cfconv racecheck 420 tests/test_gpu_cfconv.py -k "c2 or hub"
run racecheck_dense racecheck 420 tests/test_gpu_dense_tc.py -k "not large"
