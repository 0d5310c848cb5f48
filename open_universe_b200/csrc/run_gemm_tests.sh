#!/bin/bash
# Runs each gemm unit case in its own process (a trap in one must not poison the rest).
cd "$(dirname "$0")"
fail=0
for c in $(./test_gemm list); do
  timeout 120 ./test_gemm $c 0 || { fail=1; echo "--- retry $c with swapped LBO/SBO"; timeout 120 ./test_gemm $c 1; }
done
timeout 300 ./test_gemm perf
exit $fail
