#!/bin/bash
# Race / memory checking of the kernels.  ONE compute-sanitizer tool per gpurun call
# (B200_PROFILING.md), on the smallest cases: usage  scripts/gpu_sanitize.sh memcheck|racecheck|synccheck
TOOL=${1:-memcheck}
mkdir -p gpurun_out
timeout 1500 compute-sanitizer --tool $TOOL --error-exitcode 66 \
  python -m pytest tests/test_gpu_kernels.py -m gpu -q --timeout 900 -x \
  -k "comm_protocol or lr_step_gradient or kmeans_assign_simple or pagerank_fused or transitive or gram_vs_torch or monte_carlo or solve_rows" \
  > gpurun_out/sanitizer_$TOOL.log 2>&1
echo "compute-sanitizer $TOOL rc=$?"
grep -E "ERROR SUMMARY|passed|failed|Race|Invalid|Error" gpurun_out/sanitizer_$TOOL.log | tail -n 15
