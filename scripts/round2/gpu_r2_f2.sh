#!/bin/bash
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_tc_kernels.py tests/test_gpu_kernels.py -x -q --timeout 600 -k "eigh or cholesky or gram or als" > gpurun_out/f2_pytest.log 2>&1; echo "pytest rc=$?"
tail -4 gpurun_out/f2_pytest.log
timeout 120 python scripts/chol_eigh_probe.py 2>/dev/null | tail -1 | tee gpurun_out/f2_chol_eigh.json
for cfg in "0 0" "16 0" "16 16" "32 16" "8 16" "16 64"; do
  set -- $cfg
  DAPB_GRAM_PREFETCH=$1 DAPB_GRAM_PACE_EVERY=$2 DAPB_GRAM_PACE_WINDOW=$(( $2 * 2 )) timeout 300 python -m benchmarks.extras --only gram 2>gpurun_out/f2_gram.err | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('prefetch $1 pace $2:', d.get('gram_ms'), d.get('gram_hbm_gbs'), d.get('cholesky_ms'), d.get('eigh_ms'))"
done | tee gpurun_out/f2_gram_variants.txt
