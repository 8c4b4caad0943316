#!/bin/bash
cd "$(dirname "$0")/../.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -x -q --timeout 600 -k "cholesky or als" > gpurun_out/f3_pytest.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/f3_pytest.log
timeout 120 python scripts/chol_eigh_probe.py 2>/dev/null | tail -1 | tee gpurun_out/f3_chol_eigh.json
for cfg in "0 0" "16 16"; do set -- $cfg
  DAPB_GRAM_PREFETCH=$1 DAPB_GRAM_PACE_EVERY=$2 DAPB_GRAM_PACE_WINDOW=$(( $2 * 2 )) timeout 200 python scripts/gram_probe.py 2>/dev/null | tail -1
done | tee gpurun_out/f3_gram_roles.jsonl
