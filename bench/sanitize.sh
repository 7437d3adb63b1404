#!/usr/bin/env bash
# Kernel tests under compute-sanitizer on one B200 (race detection / sanitizers).
#   bash bench/sanitize.sh [memcheck|racecheck|synccheck ...]
# Small shapes only (the sanitizer slows kernels by 10-100x); TMA / tcgen05 accesses are checked by
# memcheck, shared-memory hazards of the epilogues by racecheck, barrier misuse by synccheck.
set -u
cd "$(dirname "${BASH_SOURCE[0]}")/.."
TOOLS=${*:-"memcheck racecheck"}
SEL='test_gemm_tn_plain and (128-16-64 or 256-256-32 or 333-256-24) or test_gemm_wgrad and (256-128-64 or 777-16-64) or test_gemm2_tn_pair_kernel and 256-256-64 or test_relu_bwd_bias_kernel or test_optimizer_kernels_write_bf16_shadow'
rc=0
for tool in $TOOLS; do
  echo "== compute-sanitizer --tool $tool"
  timeout 900 compute-sanitizer --tool "$tool" --error-exitcode 99 --target-processes all \
      python -m pytest tests/test_gpu_gemm.py tests/test_gpu_fused_mlp.py -m gpu -x -q -k "$SEL" 2>&1 | tail -n 15
  status=${PIPESTATUS[0]}
  [ "$status" -ne 0 ] && rc=$status
done
exit $rc
